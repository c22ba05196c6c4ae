// nodalproj.cu — single translation unit of libnodalproj.so (kernels are defined once in np_kernels.cuh).
#include "np_solver.cu"
#include "np_capi.cu"
#include "np_capi_hooks.cu"
#include "np_comm.cu"
#include "np_noalloc.cu"
#include "mac_solver.cu"
