// BEV scatter kernels (reference: networks/pointpillars8_shared.py:76-111 PointPillarsScatter.forward,
// networks/pointpillars8_export.py:44-51 Scatter.run, networks/pointpillars8_trt.py:139-205 Scatter_cuda).
//
// The canvas (B, C, nx, ny) fp32 is >= 90 % of all bytes of the front end and >= 93 % zeros.  Instead of
// memset + scattered 4-byte stores (the reference's pycuda kernel), the canvas is produced by ONE
// write-once streaming pass: a warp owns 128 consecutive cells of one frame, reads the cell -> feature-row
// map once (int4 per lane), and for every channel emits one fully coalesced 512-byte line of 16-byte
// streaming stores, pulling feature values from L2 only for the occupied lanes.  HBM-write bound.
#include "pf_internal.cuh"

namespace pf {

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

template <int MODE, bool FEAT4>
__global__ void __launch_bounds__(kFillWarps * 32) k_canvas_fill(int32_t *__restrict__ map, const float *__restrict__ feats,
                                                                 long long sv, long long sc,
                                                                 const int32_t *__restrict__ vnum, int max_voxels,
                                                                 int compact, int batch, int n_ch, int plane,
                                                                 float *__restrict__ canvas)
{
    __shared__ float s_stage[kFillWarps][kFillStage];
    __shared__ int s_rows[kFillWarps][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (plane + 127) >> 7;
    const long long wg = (long long)blockIdx.x * kFillWarps + warp;
    if (wg >= (long long)chunks * batch) return;
    const int b = (int)(wg / chunks);
    const int cell0 = (int)(wg % chunks) * 128 + lane * 4;
    long long base = (long long)b * max_voxels;
    if (MODE == 1 && compact) {   // warp-uniform: row base of frame b = voxels of the frames before it
        int s = 0;
        for (int k = lane; k < b; k += 32) s += vnum[k];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        base = s;
    }
    const bool in_plane = cell0 < plane;   // plane % 4 == 0 is guaranteed by the launcher
    int row[4] = {-1, -1, -1, -1};
    if (in_plane) {
        int32_t *mp = map + (size_t)b * plane + cell0;
        const int4 m = *reinterpret_cast<const int4 *>(mp);
        if (MODE == 0) {
            row[0] = m.x, row[1] = m.y, row[2] = m.z, row[3] = m.w;
            if ((m.x & m.y & m.z & m.w) != -1) *reinterpret_cast<int4 *>(mp) = make_int4(-1, -1, -1, -1);
        } else {
            row[0] = m.x < 0 ? (int)(base + ~m.x) : -1;
            row[1] = m.y < 0 ? (int)(base + ~m.y) : -1;
            row[2] = m.z < 0 ? (int)(base + ~m.z) : -1;
            row[3] = m.w < 0 ? (int)(base + ~m.w) : -1;
            if (m.x != kEmpty || m.y != kEmpty || m.z != kEmpty || m.w != kEmpty)
                *reinterpret_cast<int4 *>(mp) = make_int4(kEmpty, kEmpty, kEmpty, kEmpty);
        }
    }
    // slots: occupied cells of the warp, numbered in cell order
    const int mine = (row[0] >= 0) + (row[1] >= 0) + (row[2] >= 0) + (row[3] >= 0);
    int incl = mine;
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const int n_occ = __shfl_sync(0xffffffffu, incl, 31);
    float *dst = canvas + (size_t)b * n_ch * plane + cell0;
    if (n_occ == 0) {
        if (in_plane) {
#pragma unroll 8
            for (int c = 0; c < n_ch; ++c) st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
        }
        return;
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
    int cpp = ((kFillStage / n_occ) - 1) & ~3;   // channels per pass; row stride cpp+1 keeps reads conflict-free
    if (cpp > n_ch) cpp = n_ch;
    const int stride = cpp + 1;
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
            const float *p0 = stage + (slot[0] >= 0 ? slot[0] : 0) * stride;
            const float *p1 = stage + (slot[1] >= 0 ? slot[1] : 0) * stride;
            const float *p2 = stage + (slot[2] >= 0 ? slot[2] : 0) * stride;
            const float *p3 = stage + (slot[3] >= 0 ? slot[3] : 0) * stride;
#pragma unroll 4
            for (int cc = 0; cc < nc; ++cc) {
                const float a0 = slot[0] >= 0 ? p0[cc] : 0.f;
                const float a1 = slot[1] >= 0 ? p1[cc] : 0.f;
                const float a2 = slot[2] >= 0 ? p2[cc] : 0.f;
                const float a3 = slot[3] >= 0 ? p3[cc] : 0.f;
                st_cs_f4(dst + (size_t)(c0 + cc) * plane, a0, a1, a2, a3);
            }
        }
        __syncwarp();
    }
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

template <int MODE>
static cudaError_t run_fill(int32_t *map, const float *feats, int64_t sv, int64_t sc, const int32_t *vnum,
                            int max_voxels, int compact, int batch, int n_ch, int plane, float *canvas,
                            cudaStream_t stream)
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
    const long long warps = (long long)((plane + 127) / 128) * batch;
    const unsigned blocks = (unsigned)((warps + kFillWarps - 1) / kFillWarps);
    const bool feat4 = (sc == 1) && (sv % 4 == 0) && (n_ch % 4 == 0) && ((reinterpret_cast<uintptr_t>(feats) & 15) == 0);
    if (feat4)
        k_canvas_fill<MODE, true><<<blocks, kFillWarps * 32, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch,
                                                              plane, canvas);
    else
        k_canvas_fill<MODE, false><<<blocks, kFillWarps * 32, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch,
                                                               plane, canvas);
    return cudaGetLastError();
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
    const cudaError_t e = run_fill<1>(w.grid, feats, sv, sc, w.vnum, p.max_voxels, compact, n_frames, n_ch, p.plane, canvas, stream);
    mark(w, kEvFill, stream);
    return e;
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
