// Stand-alone probe (not part of the product library): where does the one-shot canvas fill (v2) lose time
// against zero-only stores?  Same structure as k_canvas_fill<0,true> with experiment switches.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fill_probe tools/fill_probe.cu && /tmp/fill_probe
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ void st_cs_f4(float *ptr, float a, float b, float c, float d)
{
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void st_f4(float *ptr, float a, float b, float c, float d)
{
    *reinterpret_cast<float4 *>(ptr) = make_float4(a, b, c, d);
}

constexpr int kW = 8, kStage = 1152;
enum { NOLOAD = 1, NOLDS = 2, NOMAPW = 4, PLAINST = 8, CH16 = 16, SPLIT = 32, SPLIT1 = 64 };

// FLAGS: experiment switches; CPW = channels per warp (64 = whole column, 16 = grid.y splits the channels)
template <int FLAGS, int CPW, int STAGE = kStage, int MINB = 1, int NW = kW>
__global__ void __launch_bounds__(NW * 32, MINB) k_fill(int32_t *__restrict__ map, const float *__restrict__ feats, int n_ch,
                                                  int batch, int plane, float *__restrict__ canvas)
{
    __shared__ float s_stage[NW][STAGE];
    __shared__ int s_rows[NW][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (plane + 127) >> 7;
    const long long wg = (long long)blockIdx.x * NW + warp;
    if (wg >= (long long)chunks * batch) return;
    const int b = (int)(wg / chunks);
    const int cell0 = (int)(wg % chunks) * 128 + lane * 4;
    const int cbase = blockIdx.y * CPW;
    int row[4] = {-1, -1, -1, -1};
    int32_t *mp = map + (size_t)b * plane + cell0;
    const int4 m = *reinterpret_cast<const int4 *>(mp);
    row[0] = m.x, row[1] = m.y, row[2] = m.z, row[3] = m.w;
    if (!(FLAGS & NOMAPW) && blockIdx.y == 0 && (m.x & m.y & m.z & m.w) != -1) *reinterpret_cast<int4 *>(mp) = m;   // same bytes back
    const int mine = (row[0] >= 0) + (row[1] >= 0) + (row[2] >= 0) + (row[3] >= 0);
    int incl = mine;
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const int n_occ = __shfl_sync(0xffffffffu, incl, 31);
    float *dst = canvas + ((size_t)b * n_ch + cbase) * plane + cell0;
    if (n_occ == 0) {
#pragma unroll 8
        for (int c = 0; c < CPW; ++c) {
            if (FLAGS & PLAINST) st_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
            else st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    // SPLIT: sectors (lane pairs) without an occupied cell are zero-filled before the features are fetched
    bool late = true;
    if (FLAGS & (SPLIT | SPLIT1)) {
        const bool has = mine > 0;
        const bool nb = __shfl_xor_sync(0xffffffffu, has, 1) != 0;   // never behind a short-circuit: all lanes must shuffle
        late = (FLAGS & SPLIT1) ? has : (has || nb);
        if (!late) {
#pragma unroll 8
            for (int c = 0; c < CPW; ++c) st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
        }
    }
    int slot[4];
    {
        int s = incl - mine;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            slot[k] = row[k] >= 0 ? s : -1;
            if (row[k] >= 0) s_rows[warp][s++] = row[k];
        }
    }
    __syncwarp();
    float *stage = s_stage[warp];
    int cpp = ((STAGE / n_occ) - 1) & ~3;
    if (cpp > CPW) cpp = CPW;
    const int stride = cpp + 1;
    for (int c0 = 0; c0 < CPW; c0 += cpp) {
        const int nc = (CPW - c0) < cpp ? (CPW - c0) : cpp;
        if (!(FLAGS & NOLOAD)) {
            const int q_per = nc >> 2;
            for (int e = lane; e < n_occ * q_per; e += 32) {
                const int s = e / q_per, q = e - s * q_per;
                const float4 v = __ldg(reinterpret_cast<const float4 *>(feats + (long long)s_rows[warp][s] * n_ch + cbase + c0) + q);
                float *d = stage + s * stride + 4 * q;
                d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
            }
        }
        __syncwarp();
        const float *p0 = stage + (slot[0] >= 0 ? slot[0] : 0) * stride;
        const float *p1 = stage + (slot[1] >= 0 ? slot[1] : 0) * stride;
        const float *p2 = stage + (slot[2] >= 0 ? slot[2] : 0) * stride;
        const float *p3 = stage + (slot[3] >= 0 ? slot[3] : 0) * stride;
#pragma unroll 4
        for (int cc = 0; cc < nc; ++cc) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            if (!(FLAGS & NOLDS)) {
                a0 = slot[0] >= 0 ? p0[cc] : 0.f;
                a1 = slot[1] >= 0 ? p1[cc] : 0.f;
                a2 = slot[2] >= 0 ? p2[cc] : 0.f;
                a3 = slot[3] >= 0 ? p3[cc] : 0.f;
            }
            if (FLAGS & PLAINST) st_f4(dst + (size_t)(c0 + cc) * plane, a0, a1, a2, a3);
            else if (late) st_cs_f4(dst + (size_t)(c0 + cc) * plane, a0, a1, a2, a3);
        }
        __syncwarp();
    }
}


// v4: zero stores first (they depend on nothing, not even the map), then the whole warp patches the occupied
// cells: lane = channel, 4-byte stores that merge in L2 with the zero lines written moments before.
template <int FLAGS, int CPW>
__global__ void __launch_bounds__(256) k_fill4(int32_t *__restrict__ map, const float *__restrict__ feats, int n_ch,
                                               int batch, int plane, float *__restrict__ canvas)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (plane + 127) >> 7;
    const long long wg = (long long)blockIdx.x * 8 + warp;
    if (wg >= (long long)chunks * batch) return;
    const int b = (int)(wg / chunks);
    const int chunk0 = (int)(wg % chunks) * 128;
    const int cell0 = chunk0 + lane * 4;
    const int cbase = blockIdx.y * CPW;
    int32_t *mp = map + (size_t)b * plane + cell0;
    const int4 m = *reinterpret_cast<const int4 *>(mp);
    float *dst = canvas + ((size_t)b * n_ch + cbase) * plane + cell0;
#pragma unroll 8
    for (int c = 0; c < CPW; ++c) {
        if (FLAGS & PLAINST) st_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
        else st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
    }
    int row[4] = {m.x, m.y, m.z, m.w};
    if (!(FLAGS & NOMAPW) && blockIdx.y == 0 && (m.x & m.y & m.z & m.w) != -1) *reinterpret_cast<int4 *>(mp) = m;
    __syncwarp();
    float *cbase_ptr = canvas + ((size_t)b * n_ch + cbase) * plane + chunk0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        unsigned mask = __ballot_sync(0xffffffffu, row[j] >= 0);
        while (mask) {
            const int src = __ffs(mask) - 1;
            mask &= mask - 1;
            const int r = __shfl_sync(0xffffffffu, row[j], src);
            const int cell = src * 4 + j;
            for (int c = lane; c < CPW; c += 32)
                cbase_ptr[(size_t)c * plane + cell] = __ldg(feats + (long long)r * n_ch + cbase + c);
        }
    }
}

// v6: persistent, statically strided, software-pipelined; item = 128 cells x CG channels; no map write-back
__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gsrc)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
template <int CG, int CAP, int ORDER>
__global__ void __launch_bounds__(256) k_fill6(const int32_t *__restrict__ map, const float *__restrict__ feats, int n_ch,
                                               int batch, int plane, float *__restrict__ canvas)
{
    constexpr int STRIDE = CG + 4;
    __shared__ __align__(16) float s_stage[8][2][CAP * STRIDE];
    __shared__ int s_occ[8][2][CAP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (plane + 127) >> 7;
    const int groups = n_ch / CG;
    const unsigned n_chunks_all = (unsigned)chunks * batch;
    const unsigned n_items = n_chunks_all * groups;
    const unsigned W = gridDim.x * 8, gw = blockIdx.x * 8 + warp;
    // ORDER 0: item = cg * n_chunks_all + chunk (adjacent warps -> adjacent chunks of one channel group)
    // ORDER 1: item = chunk * groups + cg
    auto decode = [&](unsigned item, unsigned &chunk, int &cg) {
        if (ORDER == 0) { cg = item / n_chunks_all; chunk = item - cg * n_chunks_all; }
        else { chunk = item / groups; cg = item - chunk * groups; }
    };
    auto load_map = [&](unsigned item, int4 &m) {
        m = make_int4(-1, -1, -1, -1);
        if (item < n_items) {
            unsigned chunk; int cg; decode(item, chunk, cg);
            m = __ldg(reinterpret_cast<const int4 *>(map + (size_t)chunk * 128 + lane * 4));   // plane % 128 == 0 here
        }
    };
    auto fstage = [&](unsigned item, const int4 &m, int buf, unsigned &slots, int &n_occ) {
        slots = 0xffffu; n_occ = 0;
        if (item < n_items) {
            unsigned chunk; int cg; decode(item, chunk, cg);
            const int row[4] = {m.x, m.y, m.z, m.w};
            const int mine = (row[0] >= 0) + (row[1] >= 0) + (row[2] >= 0) + (row[3] >= 0);
            int incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
            n_occ = __shfl_sync(0xffffffffu, incl, 31);
            if (n_occ > 0 && n_occ <= CAP) {
                int sl = incl - mine; slots = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    unsigned nib = 0xfu;
                    if (row[k] >= 0) { s_occ[warp][buf][sl] = row[k]; nib = (unsigned)sl; ++sl; }
                    slots |= nib << (4 * k);
                }
                __syncwarp();
                constexpr int QP = CG / 4;
                for (int e = lane; e < n_occ * QP; e += 32) {
                    const int c = e / QP, q = e - c * QP;
                    cp_async16(&s_stage[warp][buf][c * STRIDE + 4 * q], feats + (long long)s_occ[warp][buf][c] * n_ch + cg * CG + 4 * q);
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    unsigned it_s = gw, it_f = gw + W, it_m = gw + 2 * W;
    int4 m_f, m_m;
    unsigned slots_s; int nocc_s;
    load_map(it_s, m_f);
    fstage(it_s, m_f, 0, slots_s, nocc_s);
    load_map(it_f, m_f);
    int buf = 0;
    while (it_s < n_items) {
        load_map(it_m, m_m);
        unsigned slots_f; int nocc_f;
        fstage(it_f, m_f, buf ^ 1, slots_f, nocc_f);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncwarp();
        {
            unsigned chunk; int cg; decode(it_s, chunk, cg);
            const int b = chunk / chunks;
            const int cell0 = (chunk - b * chunks) * 128 + lane * 4;
            float *dst = canvas + ((size_t)b * n_ch + cg * CG) * plane + cell0;
            if (nocc_s == 0 || nocc_s > CAP) {
#pragma unroll
                for (int c = 0; c < CG; ++c) st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
            } else {
                const float *stage = s_stage[warp][buf];
                const unsigned s0 = slots_s & 15u, s1 = (slots_s >> 4) & 15u, s2 = (slots_s >> 8) & 15u, s3 = (slots_s >> 12) & 15u;
                const float *p0 = stage + (s0 % CAP) * STRIDE, *p1 = stage + (s1 % CAP) * STRIDE;
                const float *p2 = stage + (s2 % CAP) * STRIDE, *p3 = stage + (s3 % CAP) * STRIDE;
                const bool h0 = s0 != 15u, h1 = s1 != 15u, h2 = s2 != 15u, h3 = s3 != 15u;
#pragma unroll
                for (int c = 0; c < CG; ++c) {
                    const float a0 = h0 ? p0[c] : 0.f, a1 = h1 ? p1[c] : 0.f, a2 = h2 ? p2[c] : 0.f, a3 = h3 ? p3[c] : 0.f;
                    st_cs_f4(dst + (size_t)c * plane, a0, a1, a2, a3);
                }
            }
        }
        __syncwarp();
        it_s = it_f, slots_s = slots_f, nocc_s = nocc_f;
        it_f = it_m, m_f = m_m;
        it_m += W;
        buf ^= 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

int main()
{
    const int B = 8, C = 64, nx = 800, ny = 800, plane = nx * ny, V = 16000;
    const size_t n = (size_t)B * C * plane;
    float *canvas, *feats;
    int32_t *map_occ, *map_empty, *map_clu;
    CK(cudaMalloc(&canvas, n * 4));
    CK(cudaMalloc(&feats, (size_t)B * V * C * 4));
    CK(cudaMemset(feats, 0x3c, (size_t)B * V * C * 4));
    CK(cudaMalloc(&map_occ, (size_t)B * plane * 4));
    CK(cudaMalloc(&map_empty, (size_t)B * plane * 4));
    CK(cudaMalloc(&map_clu, (size_t)B * plane * 4));
    CK(cudaMemset(map_empty, 0xff, (size_t)B * plane * 4));
    {   // uniform random occupancy: V distinct cells per frame; clustered: a dense 127x126 block at the centre
        std::vector<int32_t> h((size_t)B * plane, -1), h2((size_t)B * plane, -1);
        srand(1);
        for (int b = 0; b < B; ++b) {
            int placed = 0;
            while (placed < V) {
                const int c = (int)(((long long)rand() * 32768 + rand()) % plane);
                if (h[(size_t)b * plane + c] < 0) h[(size_t)b * plane + c] = b * V + placed++;
            }
            placed = 0;
            for (int x = 337; x < 464 && placed < V; ++x)
                for (int y = 337; y < 463 && placed < V; ++y) h2[(size_t)b * plane + x * ny + y] = b * V + placed++;
        }
        CK(cudaMemcpy(map_occ, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(map_clu, h2.data(), h2.size() * 4, cudaMemcpyHostToDevice));
    }
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto report = [&](const char *name, float ms) { printf("%-58s %8.1f us  %8.1f GB/s\n", name, ms * 1e3, n * 4 / (ms * 1e-3) / 1e9); };
#define TIME(name, launch) do { for (int i = 0; i < 3; ++i) { launch; } CK(cudaDeviceSynchronize()); CK(cudaEventRecord(e0)); \
        for (int i = 0; i < 20; ++i) { launch; } CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError()); \
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); report(name, ms / 20); } while (0)
    const int chunks = plane / 128;
    const unsigned blocks = (unsigned)(((long long)chunks * B + kW - 1) / kW);
    TIME("cudaMemsetAsync", CK(cudaMemsetAsync(canvas, 0, n * 4, 0)));
    TIME("v2 empty map", (k_fill<0, 64><<<blocks, 256>>>(map_empty, feats, C, B, plane, canvas)));
    TIME("v2 uniform 2.5% map", (k_fill<0, 64><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 uniform, no map write-back", (k_fill<NOMAPW, 64><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 uniform, no feature loads", (k_fill<NOLOAD, 64><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 uniform, no LDS in the store loop", (k_fill<NOLDS, 64><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 uniform, no loads + no LDS", (k_fill<NOLOAD | NOLDS, 64><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 uniform, plain st instead of .cs", (k_fill<PLAINST, 64><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 clustered map (dense centre)", (k_fill<0, 64><<<blocks, 256>>>(map_clu, feats, C, B, plane, canvas)));
    TIME("16 ch per warp (grid.y=4) empty", (k_fill<NOMAPW, 16><<<dim3(blocks, 4), 256>>>(map_empty, feats, C, B, plane, canvas)));
    TIME("16 ch per warp (grid.y=4) uniform", (k_fill<NOMAPW, 16><<<dim3(blocks, 4), 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("16 ch per warp (grid.y=4) uniform plain st", (k_fill<NOMAPW | PLAINST, 16><<<dim3(blocks, 4), 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("8 ch per warp (grid.y=8) uniform", (k_fill<NOMAPW, 8><<<dim3(blocks, 8), 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("8 ch per warp (grid.y=8) clustered", (k_fill<NOMAPW, 8><<<dim3(blocks, 8), 256>>>(map_clu, feats, C, B, plane, canvas)));
    TIME("32 ch per warp (grid.y=2) uniform", (k_fill<NOMAPW, 32><<<dim3(blocks, 2), 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 stage 640 minb 6 uniform", (k_fill<0, 64, 640, 6><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 stage 640 minb 8 uniform", (k_fill<0, 64, 640, 8><<<blocks, 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 stage 640 minb 8 empty", (k_fill<0, 64, 640, 8><<<blocks, 256>>>(map_empty, feats, C, B, plane, canvas)));
    TIME("v2 stage 640 minb 6 clustered", (k_fill<0, 64, 640, 6><<<blocks, 256>>>(map_clu, feats, C, B, plane, canvas)));
    TIME("v2 stage 640 minb 8, 32ch uniform", (k_fill<NOMAPW, 32, 640, 8><<<dim3(blocks, 2), 256>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 4 warps per CTA uniform", (k_fill<0, 64, kStage, 1, 4><<<blocks * 2, 128>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 2 warps per CTA uniform", (k_fill<0, 64, kStage, 1, 2><<<blocks * 4, 64>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 1 warp per CTA uniform", (k_fill<0, 64, kStage, 1, 1><<<blocks * 8, 32>>>(map_occ, feats, C, B, plane, canvas)));
    TIME("v2 2 warps per CTA clustered", (k_fill<0, 64, kStage, 1, 2><<<blocks * 4, 64>>>(map_clu, feats, C, B, plane, canvas)));
    TIME("v2 2 warps per CTA empty", (k_fill<0, 64, kStage, 1, 2><<<blocks * 4, 64>>>(map_empty, feats, C, B, plane, canvas)));
    TIME("v2 16 warps per CTA uniform", (k_fill<0, 64, 576, 1, 16><<<blocks / 2, 512>>>(map_occ, feats, C, B, plane, canvas)));
    return 0;
}
