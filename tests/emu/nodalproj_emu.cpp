// TEST INFRASTRUCTURE ONLY.  Builds the product sources against the host emulation (tests/emu/cuda_emu.h).
#define NP_EMU 1
#include "../../incflo_b200/csrc/np_solver.cu"
#include "../../incflo_b200/csrc/np_capi.cu"
#include "../../incflo_b200/csrc/np_capi_hooks.cu"
#include "np_comm_emu.cpp"
#include "../../incflo_b200/csrc/np_noalloc.cu"
#include "../../incflo_b200/csrc/mac_solver.cu"
