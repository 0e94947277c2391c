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
// FEAT4: features are row-major with 16-byte aligned rows -> one float4 load covers 4 channels of a pillar.
template <int MODE, bool FEAT4>
__global__ void __launch_bounds__(256) k_canvas_fill(int32_t *__restrict__ map, const float *__restrict__ feats,
                                                     long long sv, long long sc, const int32_t *__restrict__ vnum,
                                                     int max_voxels, int compact, int batch, int n_ch, int plane,
                                                     float *__restrict__ canvas)
{
    const int lane = threadIdx.x & 31;
    const int chunks = (plane + 127) >> 7;
    const long long wg = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
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
    if (cell0 >= plane) return;   // plane % 4 == 0 is guaranteed by the launcher

    int32_t *mp = map + (size_t)b * plane + cell0;
    const int4 m = *reinterpret_cast<const int4 *>(mp);
    long long row[4];
    if (MODE == 0) {
        row[0] = m.x, row[1] = m.y, row[2] = m.z, row[3] = m.w;
        if ((m.x & m.y & m.z & m.w) != -1) *reinterpret_cast<int4 *>(mp) = make_int4(-1, -1, -1, -1);
    } else {
        row[0] = m.x < 0 ? base + ~m.x : -1;
        row[1] = m.y < 0 ? base + ~m.y : -1;
        row[2] = m.z < 0 ? base + ~m.z : -1;
        row[3] = m.w < 0 ? base + ~m.w : -1;
        if (m.x != kEmpty || m.y != kEmpty || m.z != kEmpty || m.w != kEmpty)
            *reinterpret_cast<int4 *>(mp) = make_int4(kEmpty, kEmpty, kEmpty, kEmpty);
    }
    float *dst = canvas + (size_t)b * n_ch * plane + cell0;
    const bool any = row[0] >= 0 || row[1] >= 0 || row[2] >= 0 || row[3] >= 0;
    if (!any) {
#pragma unroll 8
        for (int c = 0; c < n_ch; ++c) st_cs_f4(dst + (size_t)c * plane, 0.f, 0.f, 0.f, 0.f);
        return;
    }
    if (FEAT4) {
        for (int c = 0; c < n_ch; c += 4) {
            float4 q[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                q[k] = row[k] >= 0 ? __ldg(reinterpret_cast<const float4 *>(feats + row[k] * sv + c))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
            st_cs_f4(dst + (size_t)(c + 0) * plane, q[0].x, q[1].x, q[2].x, q[3].x);
            st_cs_f4(dst + (size_t)(c + 1) * plane, q[0].y, q[1].y, q[2].y, q[3].y);
            st_cs_f4(dst + (size_t)(c + 2) * plane, q[0].z, q[1].z, q[2].z, q[3].z);
            st_cs_f4(dst + (size_t)(c + 3) * plane, q[0].w, q[1].w, q[2].w, q[3].w);
        }
    } else {
        for (int c = 0; c < n_ch; ++c) {
            float q[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = row[k] >= 0 ? __ldg(feats + row[k] * sv + c * sc) : 0.f;
            st_cs_f4(dst + (size_t)c * plane, q[0], q[1], q[2], q[3]);
        }
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
    const unsigned blocks = (unsigned)((warps + 7) / 8);
    const bool feat4 = (sc == 1) && (sv % 4 == 0) && (n_ch % 4 == 0) && ((reinterpret_cast<uintptr_t>(feats) & 15) == 0);
    if (feat4)
        k_canvas_fill<MODE, true><<<blocks, 256, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch,
                                                              plane, canvas);
    else
        k_canvas_fill<MODE, false><<<blocks, 256, 0, stream>>>(map, feats, sv, sc, vnum, max_voxels, compact, batch, n_ch,
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
    return run_fill<1>(w.grid, feats, sv, sc, w.vnum, p.max_voxels, compact, n_frames, n_ch, p.plane, canvas, stream);
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
