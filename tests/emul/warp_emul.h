// TEST INFRASTRUCTURE ONLY.  cuda_emul.h plus the warp-level vocabulary of the voxelisation kernels
// (3d_object_detection_b200/csrc/pf_voxel_kernels.cuh): ballots, shuffles, __syncwarp, global atomics, flag polling.
// One OS thread per CUDA thread; the 32 threads of a warp meet at a per-warp barrier for every collective (every
// collective in those kernels is executed by all 32 lanes with the full mask).  The blocks of a grid run one after
// another in blockIdx order, so a block that waits for flags of blocks started EARLIER finds them set -- the same
// forward-progress argument the device code makes with its ticket counter.  Nothing in the product includes this.
#pragma once

#include "cuda_emul.h"

#include <sched.h>

struct int4 {
    int x, y, z, w;
};
struct uint4 {
    unsigned x, y, z, w;
};
static inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }

typedef void *cudaEvent_t;
typedef void *cudaStream_t;
typedef int cudaError_t;

#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __align__(n) alignas(n)

struct pf_emul_warp_ctx {
    std::mutex m;
    std::condition_variable cv;
    int waiting = 0;
    unsigned long generation = 0;
    unsigned vals[32];
    void wait()
    {
        std::unique_lock<std::mutex> lk(m);
        const unsigned long gen = generation;
        if (++waiting == 32) {
            waiting = 0, ++generation;
            cv.notify_all();
        } else
            cv.wait(lk, [&] { return generation != gen; });
    }
};
inline pf_emul_warp_ctx pf_emul_warps[32];

static inline pf_emul_warp_ctx &pf_emul_my_warp() { return pf_emul_warps[threadIdx.x >> 5]; }

static inline unsigned __ballot_sync(unsigned mask, int pred)
{
    pf_emul_warp_ctx &w = pf_emul_my_warp();
    w.vals[threadIdx.x & 31] = pred ? 1u : 0u;
    w.wait();
    unsigned r = 0;
    for (int l = 0; l < 32; ++l) r |= w.vals[l] << l;
    w.wait();
    return r & mask;
}
static inline unsigned pf_emul_exchange(unsigned v, int src)
{
    pf_emul_warp_ctx &w = pf_emul_my_warp();
    w.vals[threadIdx.x & 31] = v;
    w.wait();
    const unsigned r = (src >= 0 && src < 32) ? w.vals[src] : v;
    w.wait();
    return r;
}
static inline int __shfl_sync(unsigned, int v, int src) { return (int)pf_emul_exchange((unsigned)v, src & 31); }
static inline int __shfl_up_sync(unsigned, int v, int d) { return (int)pf_emul_exchange((unsigned)v, (int)(threadIdx.x & 31) - d); }
static inline int __shfl_xor_sync(unsigned, int v, int m) { return (int)pf_emul_exchange((unsigned)v, (int)(threadIdx.x & 31) ^ m); }
static inline void __syncwarp() { pf_emul_my_warp().wait(); }
static inline unsigned lanemask_lt() { return (1u << (threadIdx.x & 31)) - 1u; }
static inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
static inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
static inline unsigned __dp4a(unsigned a, unsigned b, unsigned c)
{
    for (int k = 0; k < 4; ++k) c += ((a >> (8 * k)) & 0xffu) * ((b >> (8 * k)) & 0xffu);
    return c;
}
static inline unsigned __vaddus4(unsigned a, unsigned b)
{
    unsigned r = 0;
    for (int k = 0; k < 4; ++k) {
        unsigned s = ((a >> (8 * k)) & 0xffu) + ((b >> (8 * k)) & 0xffu);
        r |= (s > 255u ? 255u : s) << (8 * k);
    }
    return r;
}
static inline float __fsub_rn(float a, float b) { return a - b; }       // compiled with -ffp-contract=off
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __int_as_float(int v)
{
    float f;
    memcpy(&f, &v, 4);
    return f;
}
template <class T>
static inline T __ldg(const T *p) { return *p; }
template <class T>
static inline T __ldcg(const T *p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }   // one indivisible L2 load on the device
static inline uint4 __ldcg(const uint4 *p) { return *p; }
static inline int atomicMin(int *p, int v)
{
    int old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old > v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {
    }
    return old;
}
static inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline void __nanosleep(unsigned) { sched_yield(); }
