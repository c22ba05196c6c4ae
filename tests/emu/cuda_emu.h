// TEST INFRASTRUCTURE ONLY — never part of libnodalproj.so.
//
// A tiny host emulation of the CUDA execution model, just enough to run this repository's kernels
// (incflo_b200/csrc/*.cuh) unchanged on a machine without a GPU, so that `pytest -m "not gpu"` can
// check kernel logic (index maps, tile/halo handling, pipeline schedules, the MG driver) against the
// oracle before GPU time is spent.  Blocks run one after another; the threads of a block are
// ucontext fibers that are switched at __syncthreads() / warp shuffles.  It says nothing about
// performance, races or memory alignment — compute-sanitizer and ncu on the B200 cover those.
#pragma once
#include <ucontext.h>
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))

struct uint3 { unsigned x, y, z; };
struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
struct double2 { double x, y; };
static inline double2 make_double2(double a, double b) { return double2{a, b}; }

namespace np_emu {
struct State {
    uint3 tid{0,0,0}, bid{0,0,0};
    dim3 bdim, gdim;
    // fibers
    std::vector<ucontext_t> ctx;
    std::vector<char*> stacks;
    std::vector<char> done;
    ucontext_t sched;
    int cur = -1;
    std::function<void()> body;
    bool fiber_mode = false;
    std::vector<unsigned char> dyn;
    std::vector<uint64_t> shfl;      // one slot per thread
    long long launches = 0;
};
inline State& st() { static State s; return s; }
inline void yield() {
    State& s = st();
    assert(s.fiber_mode && "kernel used a barrier/shuffle but was launched without NP_LAUNCH_SYNC");
    int me = s.cur;
    swapcontext(&s.ctx[me], &s.sched);
}
inline void trampoline() {
    State& s = st();
    s.body();
    s.done[s.cur] = 1;
    swapcontext(&s.ctx[s.cur], &s.sched);
}
inline void set_thread(int t) {
    State& s = st();
    s.cur = t;
    s.tid.x = t % s.bdim.x; s.tid.y = (t / s.bdim.x) % s.bdim.y; s.tid.z = t / (s.bdim.x * s.bdim.y);
}
// Runs the kernel body for every thread of every block.  sync=false: straight loops (the body must not
// use barriers/shuffles); sync=true: fibers.
inline void launch(dim3 grid, dim3 block, size_t smem, bool sync, std::function<void()> body) {
    State& s = st();
    s.launches++;
    s.bdim = block; s.gdim = grid; s.body = body; s.fiber_mode = sync;
    s.dyn.assign(smem + 16, 0);
    const int nt = block.x * block.y * block.z;
    if (sync) {
        const size_t STK = 256 * 1024;
        while ((int)s.stacks.size() < nt) s.stacks.push_back((char*)std::malloc(STK));
        s.ctx.resize(nt); s.done.assign(nt, 0); s.shfl.assign(nt, 0);
        for (unsigned bz = 0; bz < grid.z; ++bz) for (unsigned by = 0; by < grid.y; ++by) for (unsigned bx = 0; bx < grid.x; ++bx) {
            s.bid = uint3{bx, by, bz};
            std::fill(s.dyn.begin(), s.dyn.end(), 0);
            for (int t = 0; t < nt; ++t) {
                getcontext(&s.ctx[t]);
                s.ctx[t].uc_stack.ss_sp = s.stacks[t]; s.ctx[t].uc_stack.ss_size = STK; s.ctx[t].uc_link = &s.sched;
                makecontext(&s.ctx[t], (void (*)())trampoline, 0);
                s.done[t] = 0;
            }
            int alive = nt;
            while (alive > 0) {
                alive = 0;
                for (int t = 0; t < nt; ++t) if (!s.done[t]) {
                    set_thread(t);
                    swapcontext(&s.sched, &s.ctx[t]);
                    if (!s.done[t]) ++alive;
                }
            }
        }
    } else {
        for (unsigned bz = 0; bz < grid.z; ++bz) for (unsigned by = 0; by < grid.y; ++by) for (unsigned bx = 0; bx < grid.x; ++bx) {
            s.bid = uint3{bx, by, bz};
            for (int t = 0; t < nt; ++t) { set_thread(t); body(); }
        }
    }
    s.fiber_mode = false;
}
template <class T> inline T shfl_generic(T v, int src_lane_abs) {
    State& s = st();
    static_assert(sizeof(T) <= 8, "shuffle payload");
    uint64_t raw = 0; std::memcpy(&raw, &v, sizeof(T));
    s.shfl[s.cur] = raw;
    yield();                      // everyone has published
    uint64_t got = s.shfl[src_lane_abs];
    yield();                      // everyone has read
    T out; std::memcpy(&out, &got, sizeof(T));
    return out;
}
} // namespace np_emu

#define threadIdx (np_emu::st().tid)
#define blockIdx (np_emu::st().bid)
#define blockDim (np_emu::st().bdim)
#define gridDim (np_emu::st().gdim)
static const int warpSize = 32;
using std::min;
using std::max;

static inline void __syncthreads() { np_emu::yield(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { np_emu::yield(); }
static inline void __threadfence() {}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m) {
    int me = np_emu::st().cur; int base = me & ~31; return np_emu::shfl_generic(v, base + ((me & 31) ^ m));
}
template <class T> static inline T __shfl_down_sync(unsigned, T v, int d) {
    int me = np_emu::st().cur; int base = me & ~31; int l = (me & 31) + d; if (l > 31) l = me & 31;
    return np_emu::shfl_generic(v, base + l);
}
template <class T> static inline T __shfl_sync(unsigned, T v, int l) {
    int me = np_emu::st().cur; int base = me & ~31; return np_emu::shfl_generic(v, base + (l & 31));
}
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline double __drcp_rn(double x) { return 1.0 / x; }
static inline unsigned atomicAdd(unsigned* p, unsigned v) { unsigned o = *p; *p += v; return o; }
static inline void __threadfence_system() {}
static inline int atomicAdd(int* p, int v) { int o = *p; *p += v; return o; }
static inline double atomicAdd(double* p, double v) { double o = *p; *p += v; return o; }
static inline unsigned atomicInc(unsigned* p, unsigned lim) { unsigned o = *p; *p = (o >= lim) ? 0 : o + 1; return o; }
