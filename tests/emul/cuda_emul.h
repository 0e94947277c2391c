// TEST INFRASTRUCTURE ONLY.  A few lines of CUDA vocabulary on host threads, enough to run the kernel bodies of
// 3d_object_detection_b200/csrc/pf_post_kernels.cuh (which use threadIdx / blockIdx / __syncthreads / atomicAdd and
// nothing warp-level) inside the build container, which has no GPU.  One OS thread per CUDA thread, the blocks of a
// grid one after another; __shared__ becomes a function-local static (valid because blocks never overlap), dynamic
// shared memory a per-launch buffer.  Nothing in the product includes this file.
#pragma once

#include <math.h>
#include <stdint.h>
#include <string.h>

#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

struct pf_emul_uint3 {
    unsigned x, y, z;
};
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};

struct float4 {
    float x, y, z, w;
};
struct uchar4 {
    unsigned char x, y, z, w;
};
inline int __popc(unsigned int x) { return __builtin_popcount(x); }

inline thread_local pf_emul_uint3 threadIdx, blockIdx;
inline thread_local dim3 blockDim, gridDim;
inline void *pf_emul_dyn_smem = nullptr;

// barrier over the threads of the running block that have not returned yet (CUDA semantics: exited threads do not
// take part in later __syncthreads)
struct pf_emul_barrier {
    std::mutex m;
    std::condition_variable cv;
    int expected = 0, waiting = 0;
    unsigned long generation = 0;
    void reset(int n) { expected = n, waiting = 0; }
    void wait()
    {
        std::unique_lock<std::mutex> lk(m);
        const unsigned long gen = generation;
        if (++waiting == expected) {
            waiting = 0, ++generation;
            cv.notify_all();
        } else
            cv.wait(lk, [&] { return generation != gen; });
    }
    void drop()
    {
        std::unique_lock<std::mutex> lk(m);
        --expected;
        if (expected > 0 && waiting == expected) {
            waiting = 0, ++generation;
            cv.notify_all();
        }
    }
};
inline pf_emul_barrier pf_emul_bar;

#define __syncthreads() pf_emul_bar.wait()
#define __global__ static
#define __shared__ static
#define __launch_bounds__(...)

inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned int atomicAdd(unsigned int *p, unsigned int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }

template <class Body>
void pf_emul_launch(dim3 grid, int threads, size_t dyn_smem, Body body)
{
    std::vector<unsigned char> smem(dyn_smem + 16);
    pf_emul_dyn_smem = smem.data();
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                pf_emul_bar.reset(threads);
                std::vector<std::thread> pool;
                pool.reserve(threads);
                for (int t = 0; t < threads; ++t)
                    pool.emplace_back([=] {
                        threadIdx = {(unsigned)t, 0, 0};
                        blockIdx = {bx, by, bz};
                        blockDim = dim3((unsigned)threads);
                        gridDim = grid;
                        body();
                        pf_emul_bar.drop();
                    });
                for (auto &th : pool) th.join();
            }
    pf_emul_dyn_smem = nullptr;
}
