// BEV scatter kernels (reference: networks/pointpillars8_shared.py:76-111 PointPillarsScatter.forward,
// networks/pointpillars8_export.py:44-51 Scatter.run, networks/pointpillars8_trt.py:139-205 Scatter_cuda).
//
// The canvas (B, C, nx, ny) fp32 is >= 90 % of all bytes of the front end and >= 93 % zeros.  Instead of
// memset + scattered 4-byte stores (the reference's pycuda kernel), the canvas is produced by ONE
// write-once streaming pass: a warp owns 128 consecutive cells of one frame, reads the cell -> feature-row
// map once (int4 per lane), and for every channel emits one fully coalesced 512-byte line of 16-byte
// streaming stores, pulling feature values from L2 only for the occupied lanes.  HBM-write bound.
#include <cstdlib>
#include <mutex>

#include "pf_internal.cuh"

namespace pf {

constexpr int kFillDefault = -1;   // auto: 32-byte stores, one warp per 256 cells, or per 512 cells on canvases of >= 2^25 cells
                                   // (profiles/r2_fill_variants.md: 20 cm x 8 238 vs 239 us, nuScenes x 8 109 vs 118 us, 10 cm x 32 3281 vs 3176 us);
                                   // 16-byte stores, one warp per 128 cells, on canvases of one or two frames, where twice the
                                   // warps matter more than the wider store (20 cm x 1: 32.7 vs 45.3 us, 16 cm x 1: 46.9 vs 49.7 us)
constexpr long long kFillSmallCells = 1310720;   // 1.25 * 2^20 cells: two 20 cm frames, one 16 cm frame

// map[b][x*ny+y] = max(row) over pillars in that cell  (highest row wins == sequential index_put)
__global__ void __launch_bounds__(256) k_map_index(const int32_t *__restrict__ coors, int coors_stride,
                                                   long long n_voxels, int batch, int nx, int ny,
                                                   int32_t *__restrict__ map)
{
    const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n_voxels) return;
    const int32_t *c = coors + v * coors_stride;
    const int x = c[0], y = c[1];
    const int b = coors_stride >= 4 ? c[3] : 0;
    if (x < 0 || x >= nx || y < 0 || y >= ny || b < 0 || b >= batch) return;
    atomicMax(map + ((size_t)b * nx + x) * ny + y, (int)v);
}

// MODE 0: map holds feature rows (-1 = none), put back to -1 after reading.
// MODE 1: map is the voxeliser's grid (negative = ~voxel id, anything else = no voxel), put back to kEmpty;
//         feature row = row_base(frame) + voxel id.
// FEAT4: features are row-major with 16-byte aligned rows -> 16-byte loads of 4 channels of a pillar.
//
// A warp first stages the feature rows of its occupied cells in shared memory with ONE round of coalesced
// loads (the only exposed L2 latency), then streams the 64 channel lines out of shared memory / zeros.
// (The first version loaded per channel group inside the store loop: 16 dependent L2 round trips per warp,
// long-scoreboard bound at 3.3 TB/s -- profiles/r1_fill_v1_ncu.md.)
constexpr int kFillWarps = 8;
constexpr int kFillStage = 1152;   // floats of staging per warp: 17 occupied cells x 64 channels in one pass

// streaming (evict-first) 32-byte store, the widest the LSU has on sm_100 (SASS: STG.E.EF.256)
__device__ __forceinline__ void st_cs_f8(float *ptr, const float (&v)[8])
{
    asm volatile("st.global.cs.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "f"(v[0]), "f"(v[1]), "f"(v[2]),
                 "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}

// CPL = cells per lane: 4 -> a warp owns 128 cells and stores 16 bytes per lane and channel (STG.128),
//                       8 -> a warp owns 256 cells and stores 32 bytes per lane and channel (STG.256).
template <int MODE, bool FEAT4, int CPL, int WARPS = kFillWarps, int STAGE = kFillStage, int MINB = 1>
__global__ void __launch_bounds__(WARPS * 32, MINB) k_canvas_fill(int32_t *__restrict__ map, const float *__restrict__ feats,
                                                                 long long sv, long long sc,
                                                                 const int32_t *__restrict__ vnum, int max_voxels,
                                                                 int compact, int batch, int n_ch, int plane,
                                                                 float *__restrict__ canvas)
{
    constexpr int WCELLS = 32 * CPL;
    __shared__ float s_stage[WARPS][STAGE];
    __shared__ int s_rows[WARPS][WCELLS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (plane + WCELLS - 1) / WCELLS;
    const long long wg = (long long)blockIdx.x * WARPS + warp;
    pdl_trigger();
    const int b = (int)(wg / chunks);
    const int cell0 = (int)(wg % chunks) * WCELLS + lane * CPL;
    pdl_wait();   // map, row counts and feature rows come from the kernels in front (no-op when launched plainly)
    if (wg >= (long long)chunks * batch) return;
    long long base = (long long)b * max_voxels;
    if (MODE == 1 && compact) {   // warp-uniform: row base of frame b = voxels of the frames before it
        int s = 0;
        for (int k = lane; k < b; k += 32) s += vnum[k];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        base = s;
    }
    const bool in_plane = cell0 < plane;   // plane % CPL == 0 is guaranteed by the launcher
    int row[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) row[k] = -1;
    if (in_plane) {
        int32_t *mp = map + (size_t)b * plane + cell0;
        bool dirty = false;
#pragma unroll
        for (int h = 0; h < CPL / 4; ++h) {
            const int4 m = reinterpret_cast<const int4 *>(mp)[h];
            const int mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (MODE == 0) {
                    row[4 * h + k] = mm[k];
                    dirty |= mm[k] != -1;
                } else {
                    row[4 * h + k] = mm[k] < 0 ? (int)(base + ~mm[k]) : -1;
                    dirty |= mm[k] != kEmpty;
                }
            }
        }
        if (dirty) {
            const int e = MODE == 0 ? -1 : kEmpty;
#pragma unroll
            for (int h = 0; h < CPL / 4; ++h) reinterpret_cast<int4 *>(mp)[h] = make_int4(e, e, e, e);
        }
    }
    // slots: occupied cells of the warp, numbered in cell order
    int mine = 0;
#pragma unroll
    for (int k = 0; k < CPL; ++k) mine += row[k] >= 0;
    int incl = mine;
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const int n_occ = __shfl_sync(0xffffffffu, incl, 31);
    float *dst = canvas + (size_t)b * n_ch * plane + cell0;
    if (n_occ == 0) {
        if (in_plane) {
            const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
            for (int c = 0; c < n_ch; ++c) {
                if (CPL == 4) st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
                else {
#pragma unroll
                    for (int h = 0; h < CPL / 8; ++h) st_cs_f8(dst + (size_t)c * plane + 8 * h, z8);
                }
            }
        }
        return;
    }
    int slot[CPL];
    {
        int s = incl - mine;
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            slot[k] = row[k] >= 0 ? s : -1;
            if (row[k] >= 0) s_rows[warp][s++] = row[k];
        }
    }
    __syncwarp();
    float *stage = s_stage[warp];
    int cpp = ((STAGE / n_occ) - 1) & ~3;   // channels per pass; row stride cpp+1 keeps reads conflict-free
    int stride = cpp + 1;
    if (cpp < 4) cpp = 4, stride = 4;            // only CPL = 8 with > 230 of 256 cells occupied: tight staging, conflicts accepted
    if (cpp > n_ch) cpp = n_ch, stride = n_ch + 1;
    for (int c0 = 0; c0 < n_ch; c0 += cpp) {
        const int nc = (n_ch - c0) < cpp ? (n_ch - c0) : cpp;
        if (FEAT4) {
            const int q_per = nc >> 2;
            for (int e = lane; e < n_occ * q_per; e += 32) {
                const int s = e / q_per, q = e - s * q_per;
                const float4 v = __ldg(reinterpret_cast<const float4 *>(feats + (long long)s_rows[warp][s] * sv + c0) + q);
                float *d = stage + s * stride + 4 * q;
                d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
            }
        } else {
            for (int e = lane; e < n_occ * nc; e += 32) {
                const int s = e / nc, cc = e - s * nc;
                stage[s * stride + cc] = __ldg(feats + (long long)s_rows[warp][s] * sv + (long long)(c0 + cc) * sc);
            }
        }
        __syncwarp();
        if (in_plane) {
            const float *ps[CPL];
#pragma unroll
            for (int k = 0; k < CPL; ++k) ps[k] = stage + (slot[k] >= 0 ? slot[k] : 0) * stride;
#pragma unroll 4
            for (int cc = 0; cc < nc; ++cc) {
                float a[CPL < 8 ? 8 : CPL];
#pragma unroll
                for (int k = 0; k < CPL; ++k) a[k] = slot[k] >= 0 ? ps[k][cc] : 0.f;
                if (CPL == 4) st_cs_f4(dst + (size_t)(c0 + cc) * plane, a[0], a[1], a[2], a[3]);
                else {
#pragma unroll
                    for (int h = 0; h < CPL / 8; ++h) {
                        const float (&a8)[8] = *reinterpret_cast<const float (*)[8]>(a + 8 * h);
                        st_cs_f8(dst + (size_t)(c0 + cc) * plane + 8 * h, a8);
                    }
                }
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------
// TMA variants (Blackwell-native store path): the canvas leaves the SM as bulk asynchronous copies
// (cp.async.bulk.global.shared::cta, SASS UBLKCP.G.S) of whole channel lines instead of LSU stores.
//
// A CTA of 128 threads walks kTmaChunks consecutive chunks of CELLS cells with two [CT][CELLS (+4 pad)] fp32 tiles in
// shared memory; an "item" is one chunk x one group of CT channels.  A tile is zeroed ONCE; per item only the occupied
// columns are patched with the pillar rows, the tile goes out as CT bulk copies of CELLS * 4 bytes (issued by the lanes
// of warp 0, one bulk group per lane and item), and the columns are zeroed again when the buffer comes round.  The map
// read + feature fetch of item k+1 overlap the drain of item k; no store instruction touches the LSU.
//   <128, 64>: one item per chunk, 64 copies of 512 bytes  (first version; bound by the per-copy cost of the TMA unit)
//   <512, 16>: four items per chunk, 16 copies of 2 KB each (a quarter of the copies per byte)
constexpr int kTmaThreads = 128;
#ifndef PF_TMA_CHUNKS
#define PF_TMA_CHUNKS 8
#endif

__device__ __forceinline__ void bulk_store(float *gdst, const float *ssrc, unsigned bytes, unsigned long long policy, bool hint)
{
    const unsigned sa = (unsigned)__cvta_generic_to_shared(ssrc);
    if (hint)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst), "r"(sa),
                     "r"(bytes), "l"(policy)
                     : "memory");
    else
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(sa), "r"(bytes) : "memory");
}

template <int MODE, bool HINT, int CELLS, int CT>
__global__ void __launch_bounds__(kTmaThreads) k_canvas_fill_tma(int32_t *__restrict__ map, const float *__restrict__ feats,
                                                                 long long sv, long long sc,
                                                                 const int32_t *__restrict__ vnum, int max_voxels,
                                                                 int compact, int batch, int n_ch, int plane,
                                                                 int chunks_per_cta, float *__restrict__ canvas)
{
    constexpr int PITCH = CELLS + 4;    // floats per channel line in shared memory: 16-byte aligned rows, 4-way instead
                                        // of 32-way bank conflicts when a column is patched
    constexpr int CPT = CELLS / kTmaThreads;   // map entries per thread (1 or 4)
    extern __shared__ __align__(128) float s_tile[];   // [2][CT][PITCH]
    __shared__ int s_cell[CELLS], s_row[CELLS];        // occupied columns of the current chunk and their feature rows
    __shared__ int s_pcell[2][CELLS];                  // columns currently patched into each tile
    __shared__ int s_pn[2];
    __shared__ int s_wcnt[kTmaThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunks = (plane + CELLS - 1) / CELLS;
    const long long total = (long long)chunks * batch;
    const long long first = (long long)blockIdx.x * chunks_per_cta;
    if (first >= total) return;
    const int n_here = (int)((total - first) < chunks_per_cta ? (total - first) : chunks_per_cta);
    const int groups = (n_ch + CT - 1) / CT;
    constexpr int TILE = CT * PITCH;
    for (int j = tid * 4; j < 2 * TILE; j += kTmaThreads * 4) *reinterpret_cast<float4 *>(s_tile + j) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid < 2) s_pn[tid] = 0;
    unsigned long long policy = 0;
    if (HINT) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    __syncthreads();

    int item = 0;
    for (int k = 0; k < n_here; ++k) {
        const long long wg = first + k;
        const int b = (int)(wg / chunks);
        const int cell0 = (int)(wg % chunks) * CELLS;
        const int ncell = (plane - cell0) < CELLS ? (plane - cell0) : CELLS;   // multiple of 4 (launcher)
        // 1. map entries of my cells -> feature rows (and put the map back); compact the occupied ones
        int row[CPT];
        int mine = 0;
        {
            long long base = (long long)b * max_voxels;
            if (MODE == 1 && compact) {
                base = 0;
                for (int q = 0; q < b; ++q) base += vnum[q];
            }
            int32_t *mp = map + (size_t)b * plane + cell0 + tid * CPT;
            int m[CPT];
#pragma unroll
            for (int q = 0; q < CPT; ++q) m[q] = MODE == 0 ? -1 : kEmpty;
            if (tid * CPT < ncell) {
                if (CPT == 4) {
                    const int4 v = *reinterpret_cast<const int4 *>(mp);
                    m[0] = v.x, m[CPT > 1 ? 1 : 0] = v.y, m[CPT > 2 ? 2 : 0] = v.z, m[CPT > 3 ? 3 : 0] = v.w;
                } else {
                    m[0] = *mp;
                }
            }
            bool dirty = false;
#pragma unroll
            for (int q = 0; q < CPT; ++q) {
                if (MODE == 0) row[q] = m[q], dirty |= m[q] != -1;
                else row[q] = m[q] < 0 ? (int)(base + ~m[q]) : -1, dirty |= m[q] != kEmpty;
                mine += row[q] >= 0;
            }
            if (dirty) {
                const int e = MODE == 0 ? -1 : kEmpty;
                if (CPT == 4) *reinterpret_cast<int4 *>(mp) = make_int4(e, e, e, e);
                else *mp = e;
            }
        }
        int incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        __syncthreads();   // every thread is done with the previous chunk's lists
        if (lane == 31) s_wcnt[warp] = incl;
        __syncthreads();
        int pos = incl - mine, n_occ = 0;
#pragma unroll
        for (int q = 0; q < kTmaThreads / 32; ++q) {
            if (q < warp) pos += s_wcnt[q];
            n_occ += s_wcnt[q];
        }
#pragma unroll
        for (int q = 0; q < CPT; ++q)
            if (row[q] >= 0) s_cell[pos] = tid * CPT + q, s_row[pos] = row[q], ++pos;
        __syncthreads();

        for (int g = 0; g < groups; ++g, ++item) {
            const int buf = item & 1;
            float *tile = s_tile + buf * TILE;
            const int c0 = g * CT;
            const int nc = (n_ch - c0) < CT ? (n_ch - c0) : CT;
            // 2. the bulk copies that last read this tile (item - 2) must have finished reading shared memory
            if (warp == 0 && item >= 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncthreads();
            // 3. zero the columns patched last time, patch the new ones
            const int n_old = s_pn[buf];
            for (int e = tid; e < n_old * CT; e += kTmaThreads) {
                const int s = e / CT, c = e - s * CT;
                tile[c * PITCH + s_pcell[buf][s]] = 0.f;
            }
            __syncthreads();
            for (int e = tid; e < n_occ; e += kTmaThreads) s_pcell[buf][e] = s_cell[e];
            if (tid == 0) s_pn[buf] = n_occ;
            for (int e = tid; e < n_occ * nc; e += kTmaThreads) {
                const int s = e / nc, c = e - s * nc;
                tile[c * PITCH + s_cell[s]] = __ldg(feats + (long long)s_row[s] * sv + (long long)(c0 + c) * sc);
            }
            // 4. generic-proxy writes -> visible to the async proxy, then the bulk copies of this item
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (warp == 0) {
                float *dst = canvas + ((size_t)b * n_ch + c0) * plane + cell0;
                for (int c = lane; c < nc; c += 32) bulk_store(dst + (size_t)c * plane, tile + c * PITCH, (unsigned)ncell * 4u, policy, HINT);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (warp == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // shared memory must outlive the copies
    __syncthreads();
}

// Fallback for planes that are not a multiple of 4 cells (no shipped config): one thread per canvas element.
template <int MODE>
__global__ void __launch_bounds__(256) k_canvas_fill_scalar(int32_t *__restrict__ map, const float *__restrict__ feats,
                                                            long long sv, long long sc,
                                                            const int32_t *__restrict__ vnum, int max_voxels,
                                                            int compact, int batch, int n_ch, int plane,
                                                            float *__restrict__ canvas)
{
    const long long total = (long long)batch * n_ch * plane;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const int cell = (int)(e % plane);
        const int c = (int)((e / plane) % n_ch);
        const int b = (int)(e / ((long long)plane * n_ch));
        const int m = map[(size_t)b * plane + cell];
        long long row = -1;
        if (MODE == 0) {
            row = m;
        } else if (m < 0) {
            long long base = (long long)b * max_voxels;
            if (compact) { base = 0; for (int k = 0; k < b; ++k) base += vnum[k]; }
            row = base + ~m;
        }
        canvas[e] = row >= 0 ? feats[row * sv + c * sc] : 0.f;
    }
}

__global__ void __launch_bounds__(256) k_map_reset(int32_t *__restrict__ map, size_t n, int32_t value)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        map[i] = value;
}

// grad_feats(v, c) = grad_canvas[b][c][x*ny + y]  (autograd of the index_put)
__global__ void __launch_bounds__(256) k_scatter_backward(const float *__restrict__ grad_canvas,
                                                          const int32_t *__restrict__ coors, int coors_stride,
                                                          long long n_voxels, int batch, int n_ch, int nx, int ny,
                                                          float *__restrict__ grad_feats, long long sv, long long sc)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_voxels * n_ch) return;
    const long long v = e / n_ch;
    const int c = (int)(e % n_ch);
    const int32_t *co = coors + v * coors_stride;
    const int x = co[0], y = co[1];
    const int b = coors_stride >= 4 ? co[3] : 0;
    float g = 0.f;
    if (x >= 0 && x < nx && y >= 0 && y < ny && b >= 0 && b < batch)
        g = grad_canvas[(((size_t)b * n_ch + c) * nx + x) * ny + y];
    grad_feats[v * sv + c * sc] = g;
}

// fill variants (pf_debug_set_fill_variant / PF_FILL_VARIANT): 0 = warp, 16-byte stores (round 1's kernel), 1 = warp,
// 32-byte stores, 2 / 3 = TMA bulk stores of 512 bytes (3: L2 evict-first hint), 4 / 5 = TMA bulk stores of 2 KB
static int g_fill_variant = -1;
static bool g_fill_set = false;

int fill_variant()
{
    if (!g_fill_set) {
        const char *e = getenv("PF_FILL_VARIANT");
        g_fill_variant = (e && e[0] >= '0' && e[0] <= '8') ? e[0] - '0' : kFillDefault;
        g_fill_set = true;
    }
    return g_fill_variant;
}
void set_fill_variant(int v) { g_fill_variant = (v >= 0 && v <= 8) ? v : kFillDefault, g_fill_set = true; }

template <int MODE>
static cudaError_t run_fill(int32_t *map, const float *feats, int64_t sv, int64_t sc, const int32_t *vnum,
                            int max_voxels, int compact, int batch, int n_ch, int plane, float *canvas,
                            cudaStream_t stream, bool dependent = false)
{
    if (batch <= 0 || n_ch <= 0 || plane <= 0) return cudaSuccess;
    const bool vec_ok = (plane % 4 == 0) && ((reinterpret_cast<uintptr_t>(canvas) & 15) == 0);
    if (!vec_ok) {
        k_canvas_fill_scalar<MODE><<<148 * 8, 256, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch,
                                                              n_ch, plane, canvas);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        k_map_reset<<<148 * 4, 256, 0, stream>>>(map, (size_t)batch * plane, MODE == 0 ? -1 : kEmpty);
        return cudaGetLastError();
    }
    int variant = fill_variant();
    if (variant < 0) {   // the rule behind kFillDefault
        const long long cells = (long long)batch * plane;
        variant = cells >= (1ll << 25) ? 6 : (cells <= kFillSmallCells ? 0 : 1);
    }
    const bool feat4 = (sc == 1) && (sv % 4 == 0) && (n_ch % 4 == 0) && ((reinterpret_cast<uintptr_t>(feats) & 15) == 0);
    if (variant >= 2 && variant <= 5) {
        // 2: 128-cell chunks x 64 channels (512-byte copies); 3: the same + L2 evict-first; 4: 512-cell chunks x 16 channels
        // (2 KB copies); 5: 4 + evict-first
        const bool big = variant >= 4;
        const int cells = big ? 512 : 128, ct = big ? 16 : 64;
        if (big || n_ch <= 64) {
            const long long chunks = (long long)((plane + cells - 1) / cells) * batch;
            const int per_cta = big ? (PF_TMA_CHUNKS / 4 > 0 ? PF_TMA_CHUNKS / 4 : 1) : PF_TMA_CHUNKS;
            const unsigned blocks = (unsigned)((chunks + per_cta - 1) / per_cta);
            const size_t smem = (size_t)2 * ct * (cells + 4) * sizeof(float);
            void (*kern)(int32_t *, const float *, long long, long long, const int32_t *, int, int, int, int, int, int, float *);
            switch (variant) {
            case 2: kern = k_canvas_fill_tma<MODE, false, 128, 64>; break;
            case 3: kern = k_canvas_fill_tma<MODE, true, 128, 64>; break;
            case 4: kern = k_canvas_fill_tma<MODE, false, 512, 16>; break;
            default: kern = k_canvas_fill_tma<MODE, true, 512, 16>; break;
            }
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
            kern<<<blocks, kTmaThreads, smem, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, per_cta, canvas);
            return cudaGetLastError();
        }
    }
    const bool al32 = (reinterpret_cast<uintptr_t>(canvas) & 31) == 0;
    if (variant == 6 && (plane % 16 != 0 || !al32)) variant = 1;
    if (variant == 6) {   // 512 cells per warp, 4 warps per CTA, two 32-byte stores per lane and channel
        const long long warps6 = (long long)((plane + 511) / 512) * batch;
        const unsigned blocks6 = (unsigned)((warps6 + 3) / 4);
        if (feat4)
            return launch_k(k_canvas_fill<MODE, true, 16, 4, 2304>, dim3(blocks6), dim3(128), 0, stream, dependent, map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas);
        return launch_k(k_canvas_fill<MODE, false, 16, 4, 2304>, dim3(blocks6), dim3(128), 0, stream, dependent, map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas);
    }
    if ((variant == 7 || variant == 8) && (plane % 8 != 0 || !al32)) variant = 0;
    if (variant == 7 || variant == 8) {   // experiments on the 256-cell warp: 7 = 4 warps per CTA, 8 = registers capped for 5 CTAs per SM
        const long long w78 = (long long)((plane + 255) / 256) * batch;
        if (variant == 7) {
            const unsigned b7 = (unsigned)((w78 + 3) / 4);
            if (feat4) k_canvas_fill<MODE, true, 8, 4, 1152><<<b7, 128, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas);
            else k_canvas_fill<MODE, false, 8, 4, 1152><<<b7, 128, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas);
        } else {
            const unsigned b8 = (unsigned)((w78 + 7) / 8);
            if (feat4) k_canvas_fill<MODE, true, 8, 8, 1024, 5><<<b8, 256, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas);
            else k_canvas_fill<MODE, false, 8, 8, 1024, 5><<<b8, 256, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas);
        }
        return cudaGetLastError();
    }
    const bool wide = variant == 1 && (plane % 8 == 0) && al32;   // else variant 0
    const int wcells = wide ? 256 : 128;
    const long long warps = (long long)((plane + wcells - 1) / wcells) * batch;
    const unsigned blocks = (unsigned)((warps + kFillWarps - 1) / kFillWarps);
#define PF_FILL_LAUNCH(F4, CPL) launch_k(k_canvas_fill<MODE, F4, CPL>, dim3(blocks), dim3(kFillWarps * 32), 0, stream, dependent, map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch, plane, canvas)
    cudaError_t le;
    if (wide) {
        le = feat4 ? PF_FILL_LAUNCH(true, 8) : PF_FILL_LAUNCH(false, 8);
    } else {
        le = feat4 ? PF_FILL_LAUNCH(true, 4) : PF_FILL_LAUNCH(false, 4);
    }
#undef PF_FILL_LAUNCH
    return le;
}

cudaError_t launch_scatter(const Params &p, const Work &w, const float *feats, int64_t sv, int64_t sc,
                           const int32_t *coors, int coors_stride, int64_t n_voxels, int batch, int n_ch,
                           float *canvas, cudaStream_t stream)
{
    if (n_voxels > 0) {
        k_map_index<<<(unsigned)((n_voxels + 255) / 256), 256, 0, stream>>>(coors, coors_stride, n_voxels, batch, p.nx,
                                                                          p.ny, w.map);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return run_fill<0>(w.map, feats, sv, sc, nullptr, 0, 0, batch, n_ch, p.plane, canvas, stream);
}

cudaError_t launch_scatter_from_map(const Params &p, const Work &w, const float *feats, int64_t sv, int64_t sc,
                                    int n_frames, int compact, int n_ch, float *canvas, cudaStream_t stream)
{
    mark(w, kEvFillStart, stream);
    // as a programmatic dependent of whatever kernel is in front of it on the stream (the gather, or the caller's feature
    // net): launch and index arithmetic overlap that kernel's tail, the kernel waits before it reads anything
    const cudaError_t e = run_fill<1>(w.grid, feats, sv, sc, w.vnum, p.max_voxels, compact, n_frames, n_ch, p.plane, canvas, stream,
                                      pdl_stage(6) && !w.ev);
    mark(w, kEvFill, stream);
    return e;
}

// ---- allocation-free scatter (pf_scatter_static; no counterpart in the reference, added
// for callers that may not allocate: code under graph capture, tight serving loops) ----
// The cell -> feature-row map of launch_scatter lives in the handle's workspace (cudaMalloc in pf_create).  This variant
// keeps it in a file-scope device array instead: 2^22 int32 cells = 16 MB, one frame of the 10 cm grid (1600 x 1600) or
// six of the 20 cm grid.  Every GPU that loads the library holds the array until exit; every call shares it, so calls that
// overlap on the GPU (two streams, two host threads) race.
constexpr long long kStaticMapCells = PF_SCATTER_STATIC_MAX_CELLS;   // the limit stated in include/pillarfront.h
static __device__ __align__(256) int32_t g_static_map[kStaticMapCells];

// In host code `g_static_map` is a host-side stand-in, not a device address: the address the kernels get is asked from
// the runtime, once per device, with that device current.
constexpr int kStaticMaxDevices = 64;
static int32_t *g_static_map_ptr[kStaticMaxDevices] = {};
static std::mutex g_static_mutex;

static cudaError_t static_map_address(int32_t **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_static_mutex);
    if (!g_static_map_ptr[dev]) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_static_map);
        if (e != cudaSuccess) return e;
        if (!ptr) return cudaErrorInvalidSymbol;
        g_static_map_ptr[dev] = static_cast<int32_t *>(ptr);
    }
    *out = g_static_map_ptr[dev];
    return cudaSuccess;
}

long long static_scatter_max_cells() { return kStaticMapCells; }

cudaError_t launch_scatter_static(const float *feats, int64_t sv, int64_t sc, const int32_t *coors, int coors_stride,
                                  int64_t n_voxels, int batch, int n_ch, int nx, int ny, float *canvas,
                                  cudaStream_t stream)
{
    const long long plane = (long long)nx * ny;
    if (batch < 1 || nx < 1 || ny < 1 || plane * batch > kStaticMapCells) return cudaErrorInvalidValue;
    int32_t *map = nullptr;
    cudaError_t e = static_map_address(&map);
    if (e != cudaSuccess) return e;
    // the array keeps whatever the last call (of any shape, possibly one that failed half-way) left: start from "no row"
    const size_t cells = (size_t)batch * plane;
    const unsigned reset_blocks = (unsigned)((cells + 1023) / 1024 < 148 * 4 ? (cells + 1023) / 1024 : 148 * 4);
    k_map_reset<<<reset_blocks, 256, 0, stream>>>(map, cells, -1);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (n_voxels > 0) {
        k_map_index<<<(unsigned)((n_voxels + 255) / 256), 256, 0, stream>>>(coors, coors_stride, n_voxels, batch, nx, ny, map);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return run_fill<0>(map, feats, sv, sc, nullptr, 0, 0, batch, n_ch, (int)plane, canvas, stream);
}

cudaError_t launch_scatter_backward(const float *grad_canvas, const int32_t *coors, int coors_stride,
                                    int64_t n_voxels, int batch, int n_ch, int nx, int ny, float *grad_feats,
                                    int64_t sv, int64_t sc, cudaStream_t stream)
{
    if (n_voxels <= 0 || n_ch <= 0) return cudaSuccess;
    const long long total = (long long)n_voxels * n_ch;
    k_scatter_backward<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(grad_canvas, coors, coors_stride, n_voxels,
                                                                           batch, n_ch, nx, ny, grad_feats, sv, sc);
    return cudaGetLastError();
}

}  // namespace pf
