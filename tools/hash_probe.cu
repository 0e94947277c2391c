// Stand-alone probe (not part of the product library): "first point index per cell" on a dense grid against an
// open-addressing hash, the two options BASELINE.json names for the 10 cm and nuScenes grids.  Same synthetic cell
// stream for both (S1-like: K distinct cells with power-law weights), one launch over B frames.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/hash_probe tools/hash_probe.cu && /tmp/hash_probe
//
// dense : atomicMin(grid[f][cell], i) behind an ld.cg probe (what k_quantize does); the grid is never cleared by a
//         memset in the product (self-cleaning), so the timed loop restores only the touched cells.
// hash  : table[f][h] = (cell << 32 | i) as one 64-bit word, empty = ~0; insert = linear probing with a 64-bit
//         atomicMin when the slot already holds this cell, atomicCAS on an empty slot.  Load factor K / capacity.
//         The hash variant must also be cleared (or self-cleaned) per frame: capacity * 8 bytes.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__global__ void __launch_bounds__(256) k_dense(const int32_t *__restrict__ cell, int n, int cells, int32_t *__restrict__ grid)
{
    const int f = blockIdx.y;
    const int32_t *c = cell + (size_t)f * n;
    int32_t *g = grid + (size_t)f * cells;
    for (int k = 0; k < 4; ++k) {
        const int i = blockIdx.x * 1024 + k * 256 + threadIdx.x;
        if (i >= n) return;
        const int cc = c[i];
        if (cc < 0) continue;
        if (__ldcg(g + cc) > i) atomicMin(g + cc, i);
    }
}

__global__ void __launch_bounds__(256) k_dense_reset(const int32_t *__restrict__ cell, int n, int cells, int32_t *__restrict__ grid)
{
    const int f = blockIdx.y;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int cc = cell[(size_t)f * n + i];
    if (cc >= 0) grid[(size_t)f * cells + cc] = 0x7fffffff;
}

__device__ __forceinline__ unsigned hash_cell(unsigned c) { return c * 2654435761u; }

__global__ void __launch_bounds__(256) k_hash(const int32_t *__restrict__ cell, int n, int cap_log2,
                                              unsigned long long *__restrict__ table)
{
    const int f = blockIdx.y;
    const int32_t *c = cell + (size_t)f * n;
    unsigned long long *t = table + ((size_t)f << cap_log2);
    const unsigned mask = (1u << cap_log2) - 1;
    for (int k = 0; k < 4; ++k) {
        const int i = blockIdx.x * 1024 + k * 256 + threadIdx.x;
        if (i >= n) return;
        const int cc = c[i];
        if (cc < 0) continue;
        const unsigned long long mine = ((unsigned long long)(unsigned)cc << 32) | (unsigned)i;
        unsigned h = (hash_cell((unsigned)cc) >> (32 - cap_log2)) & mask;
        while (true) {
            unsigned long long cur = __ldcg(t + h);
            if (cur == ~0ull) {
                const unsigned long long old = atomicCAS(t + h, ~0ull, mine);
                if (old == ~0ull) break;
                cur = old;
            }
            if ((unsigned)(cur >> 32) == (unsigned)cc) {
                if (cur > mine) atomicMin(t + h, mine);
                break;
            }
            h = (h + 1) & mask;
        }
    }
}

int main()
{
    struct Case { const char *name; int n, cells, K, B; } cases[] = {
        {"20 cm   (800x800,   N=120k, K=12000, B=8)", 120000, 640000, 12000, 8},
        {"10 cm   (1600x1600, N=300k, K=15000, B=32)", 300000, 2560000, 15000, 32},
        {"nuScenes(512x480,   N=350k, K=9000,  B=8)", 350000, 245760, 9000, 8},
    };
    for (const Case &cs : cases) {
        // S1-like stream: K distinct cells, weights ~ k^-0.8
        std::vector<int32_t> h((size_t)cs.B * cs.n);
        std::vector<double> cdf(cs.K);
        double acc = 0;
        for (int k = 0; k < cs.K; ++k) { acc += 1.0 / pow((double)(k + 1), 0.8); cdf[k] = acc; }
        srand(7);
        for (int f = 0; f < cs.B; ++f) {
            std::vector<int32_t> ids(cs.K);
            for (int k = 0; k < cs.K; ++k) ids[k] = (int32_t)(((unsigned long long)rand() * 32768ull + rand()) % cs.cells);
            for (int i = 0; i < cs.n; ++i) {
                const double u = (double)rand() / RAND_MAX * acc;
                const int k = (int)(std::lower_bound(cdf.begin(), cdf.end(), u) - cdf.begin());
                h[(size_t)f * cs.n + i] = ids[k < cs.K ? k : cs.K - 1];
            }
        }
        int32_t *d_cell, *d_grid;
        unsigned long long *d_tab;
        int cap_log2 = 1;
        while ((1 << cap_log2) < 4 * cs.K) ++cap_log2;      // load factor <= 0.25
        CK(cudaMalloc(&d_cell, h.size() * 4));
        CK(cudaMalloc(&d_grid, (size_t)cs.B * cs.cells * 4));
        CK(cudaMalloc(&d_tab, ((size_t)cs.B << cap_log2) * 8));
        CK(cudaMemcpy(d_cell, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemset(d_grid, 0x7f, (size_t)cs.B * cs.cells * 4));
        const dim3 g4((cs.n + 1023) / 1024, cs.B), g1((cs.n + 255) / 256, cs.B);
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        float t_dense = 0, t_reset = 0, t_hash = 0, t_clear = 0, ms;
        const int reps = 20;
        for (int r = -3; r < reps; ++r) {
            CK(cudaEventRecord(e0));
            k_dense<<<g4, 256>>>(d_cell, cs.n, cs.cells, d_grid);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r >= 0) t_dense += ms;
            CK(cudaEventRecord(e0));
            k_dense_reset<<<g1, 256>>>(d_cell, cs.n, cs.cells, d_grid);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r >= 0) t_reset += ms;
            CK(cudaEventRecord(e0));
            CK(cudaMemsetAsync(d_tab, 0xff, ((size_t)cs.B << cap_log2) * 8));
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r >= 0) t_clear += ms;
            CK(cudaEventRecord(e0));
            k_hash<<<g4, 256>>>(d_cell, cs.n, cap_log2, d_tab);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r >= 0) t_hash += ms;
        }
        CK(cudaGetLastError());
        printf("%s  dense atomicMin %7.1f us (+ touched-cell reset %6.1f us, folded into k_slot / the fill in the product)"
               "   hash insert %7.1f us (+ table clear %5.1f us, capacity 2^%d per frame)\n",
               cs.name, 1e3 * t_dense / reps, 1e3 * t_reset / reps, 1e3 * t_hash / reps, 1e3 * t_clear / reps, cap_log2);
        cudaFree(d_cell), cudaFree(d_grid), cudaFree(d_tab);
    }
    return 0;
}
