// Stand-alone probe (not part of the product library): which store pattern writes a (B,64,800,800) fp32
// canvas fastest on this B200?  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/write_probe tools/write_probe.cu && /tmp/write_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <int ST> __device__ __forceinline__ void st4(float *p, float4 v)
{
    if (ST == 0) asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else if (ST == 1) *reinterpret_cast<float4 *>(p) = v;
    else if (ST == 2) asm volatile("st.global.wt.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// P1: linear grid-stride
template <int ST> __global__ void __launch_bounds__(256) k_linear(float *c, size_t n4)
{
    const float4 z = make_float4(0, 0, 0, 0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) st4<ST>(c + 4 * i, z);
}

// P2: warp = CELLS_PER_LANE*128 consecutive cells x CH channels (channel group = blockIdx.y)
template <int ST, int VPL> __global__ void __launch_bounds__(256) k_warp_cols(float *c, int batch, int n_ch, int ch_per, int plane)
{
    const int lane = threadIdx.x & 31;
    const int chunk_cells = 128 * VPL;
    const int chunks = plane / chunk_cells;
    const long long wg = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (wg >= (long long)chunks * batch) return;
    const int b = (int)(wg / chunks);
    const int cell0 = (int)(wg % chunks) * chunk_cells + lane * 4;
    const int c0 = blockIdx.y * ch_per;
    float *dst = c + ((size_t)b * n_ch + c0) * plane + cell0;
    const float4 z = make_float4(0, 0, 0, 0);
#pragma unroll 4
    for (int ch = 0; ch < ch_per; ++ch)
#pragma unroll
        for (int v = 0; v < VPL; ++v) st4<ST>(dst + (size_t)ch * plane + v * 128, z);
}

// P3: CTA = 1024*VPL cells, all warps walk the channels in lock step (barrier every SYNC channels)
template <int ST, int VPL, int SYNC> __global__ void __launch_bounds__(256) k_cta_lockstep(float *c, int batch, int n_ch, int plane)
{
    const int chunk_cells = 1024 * VPL;
    const int chunks = plane / chunk_cells;
    const int b = blockIdx.x / chunks;
    const int cell0 = (blockIdx.x % chunks) * chunk_cells + threadIdx.x * 4;
    float *dst = c + (size_t)b * n_ch * plane + cell0;
    const float4 z = make_float4(0, 0, 0, 0);
    for (int ch = 0; ch < n_ch; ++ch) {
#pragma unroll
        for (int v = 0; v < VPL; ++v) st4<ST>(dst + (size_t)ch * plane + v * 1024, z);
        if (SYNC && (ch % SYNC) == SYNC - 1) __syncthreads();
    }
}

// P4: TMA-style bulk stores (cp.async.bulk.global.shared::cta) from a zeroed shared tile
template <int BYTES> __global__ void __launch_bounds__(128) k_bulk(float *c, int batch, int n_ch, int plane, int inflight)
{
    extern __shared__ __align__(128) unsigned char tile[];
    for (int i = threadIdx.x * 16; i < BYTES; i += blockDim.x * 16) *reinterpret_cast<uint4 *>(tile + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int chunk_cells = BYTES / 4;
    const int chunks = plane / chunk_cells;
    const int b = blockIdx.x / chunks;
    const int cell0 = (blockIdx.x % chunks) * chunk_cells;
    if (threadIdx.x == 0) {
        const uint32_t s = (uint32_t)__cvta_generic_to_shared(tile);
        for (int ch = 0; ch < n_ch; ++ch) {
            float *dst = c + ((size_t)b * n_ch + ch) * plane + cell0;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(s), "r"(BYTES) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    (void)inflight;
}

int main()
{
    const int B = 8, C = 64, plane = 800 * 800;
    const size_t n = (size_t)B * C * plane;
    float *canvas;
    CK(cudaMalloc(&canvas, n * 4));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto report = [&](const char *name, float ms) { printf("%-44s %8.1f us  %8.1f GB/s\n", name, ms * 1e3, n * 4 / (ms * 1e-3) / 1e9); };
#define TIME(name, launch) do { for (int i = 0; i < 3; ++i) { launch; } CK(cudaDeviceSynchronize()); CK(cudaEventRecord(e0)); \
        for (int i = 0; i < 20; ++i) { launch; } CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError()); \
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); report(name, ms / 20); } while (0)

    TIME("cudaMemsetAsync", CK(cudaMemsetAsync(canvas, 0, n * 4, 0)));
    TIME("linear .cs   grid 148*8", (k_linear<0><<<148 * 8, 256>>>(canvas, n / 4)));
    TIME("linear st    grid 148*8", (k_linear<1><<<148 * 8, 256>>>(canvas, n / 4)));
    TIME("linear st    grid 148*32", (k_linear<1><<<148 * 32, 256>>>(canvas, n / 4)));
    TIME("linear .wt   grid 148*8", (k_linear<2><<<148 * 8, 256>>>(canvas, n / 4)));
    TIME("linear noalloc grid 148*8", (k_linear<3><<<148 * 8, 256>>>(canvas, n / 4)));
    {
        const int chunks = plane / 128; const unsigned blocks = (unsigned)(((long long)chunks * B + 7) / 8);
        TIME("warp 128 cells x 64 ch .cs (current)", (k_warp_cols<0, 1><<<dim3(blocks, 1), 256>>>(canvas, B, C, 64, plane)));
        TIME("warp 128 cells x 64 ch st", (k_warp_cols<1, 1><<<dim3(blocks, 1), 256>>>(canvas, B, C, 64, plane)));
        TIME("warp 128 cells x 64 ch .wt", (k_warp_cols<2, 1><<<dim3(blocks, 1), 256>>>(canvas, B, C, 64, plane)));
        TIME("warp 128 cells x 16 ch (x4 grid) st", (k_warp_cols<1, 1><<<dim3(blocks, 4), 256>>>(canvas, B, C, 16, plane)));
        TIME("warp 128 cells x 8 ch (x8 grid) st", (k_warp_cols<1, 1><<<dim3(blocks, 8), 256>>>(canvas, B, C, 8, plane)));
        TIME("warp 128 cells x 8 ch (x8 grid) .cs", (k_warp_cols<0, 1><<<dim3(blocks, 8), 256>>>(canvas, B, C, 8, plane)));
    }
    {
        const int chunks = plane / 512; const unsigned blocks = (unsigned)(((long long)chunks * B + 7) / 8);
        TIME("warp 512 cells x 64 ch st", (k_warp_cols<1, 4><<<dim3(blocks, 1), 256>>>(canvas, B, C, 64, plane)));
        TIME("warp 512 cells x 64 ch .cs", (k_warp_cols<0, 4><<<dim3(blocks, 1), 256>>>(canvas, B, C, 64, plane)));
        TIME("warp 512 cells x 16 ch st", (k_warp_cols<1, 4><<<dim3(blocks, 4), 256>>>(canvas, B, C, 16, plane)));
    }
    TIME("cta 1024 cells lockstep sync1 st", (k_cta_lockstep<1, 1, 1><<<plane / 1024 * B, 256>>>(canvas, B, C, plane)));
    TIME("cta 1024 cells lockstep sync8 st", (k_cta_lockstep<1, 1, 8><<<plane / 1024 * B, 256>>>(canvas, B, C, plane)));
    TIME("cta 1024 cells free-running st", (k_cta_lockstep<1, 1, 0><<<plane / 1024 * B, 256>>>(canvas, B, C, plane)));
    TIME("cta 4096 cells lockstep sync1 st", (k_cta_lockstep<1, 4, 1><<<plane / 4096 * B, 256>>>(canvas, B, C, plane)));
    TIME("cta 4096 cells lockstep sync1 .cs", (k_cta_lockstep<0, 4, 1><<<plane / 4096 * B, 256>>>(canvas, B, C, plane)));
    TIME("cta 4096 cells free-running st", (k_cta_lockstep<1, 4, 0><<<plane / 4096 * B, 256>>>(canvas, B, C, plane)));
    TIME("bulk 4 KB tile  (1024 cells)", (k_bulk<4096><<<plane / 1024 * B, 128, 4096>>>(canvas, B, C, plane, 0)));
    TIME("bulk 16 KB tile (4096 cells)", (k_bulk<16384><<<plane / 4096 * B, 128, 16384>>>(canvas, B, C, plane, 0)));
    TIME("bulk 32 KB tile (6400 cells)", (k_bulk<25600><<<plane / 6400 * B, 128, 25600>>>(canvas, B, C, plane, 0)));
    cudaFree(canvas);
    return 0;
}
