// Anchor occupancy mask ("next" row N2): which anchors have at least one occupied pillar under their bird's-eye
// rectangle.  Replaces fused_get_anchors_mask_gpu of the reference (framework/box_np_ops.py:236-257 = init_map_gpu
// :168-172, cumx_gpu :175-180, cumy_gpu :183-188, get_anchors_mask_gpu :214-223), called per frame from
// AnchorAssigner.create_mask (framework/anchor_assigner.py:322-335).
//
// Same arithmetic, different schedule.  The reference uploads a zeroed (nx, ny) fp32 map per frame, adds 1 per
// pillar, runs two kernels whose threads each walk a whole row / column serially, and looks four corners up per
// anchor.  Here the map is an integer table (the counts are integers <= max_voxels < 2^24, so the reference's fp32
// sums are exact and an integer table gives the same comparison), a batch of frames is processed by every launch,
// the two scans are parallel (warp per row with shuffle scans; 32-column strips with a two-pass block scan), and
// one thread per anchor reads its rectangle once and tests it against up to eight frames.  The lookup is bound by
// L2 sectors (four corner reads per anchor and frame), so the table is uint16 whenever a frame cannot hold more
// than 65535 pillars (every shipped config: max_voxels 12000-20000) and int32 otherwise; the tables of a frame
// group stay in the 126 MB L2 between the four kernels (8 frames x 1.28 MB at 20 cm).
#include "pf_internal.cuh"

namespace pf {

namespace {

constexpr int kAmThreads = 256;
constexpr int kAmFrameGroup = 8;   // frames one lookup thread walks; their tables share the L2

// rows: coors[r*stride + {0,1}] = (x, y); frame of row r:
//   voxel_num == nullptr            : every row valid, frame = coors[r*stride+3] when n_frames > 1 (merge_second_batch
//                                     layout, framework/utils.py:38-41), else 0
//   voxel_num, rows_per_frame  > 0  : slab layout of pf_voxelize (frame f owns rows [f*rows_per_frame, +voxel_num[f]))
//   voxel_num, rows_per_frame == 0  : compact layout of pf_voxelize (first sum(voxel_num) rows, frame id in column 3)
__device__ __forceinline__ void table_add_one(int32_t *t, size_t i) { atomicAdd(t + i, 1); }
// uint16 cells: add into the containing 32-bit word; a cell never exceeds 65535, so no carry crosses the halves
__device__ __forceinline__ void table_add_one(uint16_t *t, size_t i)
{
    atomicAdd(reinterpret_cast<unsigned int *>(t) + (i >> 1), (i & 1) ? 0x10000u : 1u);
}

template <typename T>
__global__ void __launch_bounds__(kAmThreads)
k_am_count(const int32_t *__restrict__ coors, int stride, const int32_t *__restrict__ voxel_num, long long n_rows,
           int rows_per_frame, int n_frames, int nx, int ny, T *__restrict__ integ)
{
    const long long r = (long long)blockIdx.x * kAmThreads + threadIdx.x;
    if (r >= n_rows) return;
    int f = 0;
    bool frame_from_column = n_frames > 1;
    if (voxel_num) {
        if (rows_per_frame > 0) {
            f = (int)(r / rows_per_frame);
            if (f >= n_frames || r - (long long)f * rows_per_frame >= voxel_num[f]) return;
            frame_from_column = false;
        } else {
            long long total = 0;
            for (int k = 0; k < n_frames; ++k) total += voxel_num[k];
            if (r >= total) return;
        }
    }
    const int32_t *c = coors + r * stride;
    if (frame_from_column) f = c[3];
    const int x = c[0], y = c[1];
    if ((unsigned)f >= (unsigned)n_frames || (unsigned)x >= (unsigned)nx || (unsigned)y >= (unsigned)ny) return;
    table_add_one(integ, ((size_t)f * nx + x) * ny + y);
}

__device__ __forceinline__ int warp_inclusive_scan(int v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// cumy_gpu: running sum along y (the contiguous axis); one warp per (frame, x) row, 128 cells per trip.
template <typename T>
__global__ void __launch_bounds__(kAmThreads)
k_am_scan_y(T *__restrict__ integ, long long n_rows, int ny)
{
    const long long row = ((long long)blockIdx.x * kAmThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= n_rows) return;
    T *p = integ + row * ny;
    int carry = 0;
    for (int base = 0; base < ny; base += 128) {
        int v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = base + k * 32 + lane;
            v[k] = i < ny ? p[i] : 0;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v[k] = warp_inclusive_scan(v[k], lane) + carry;
            carry = __shfl_sync(0xffffffffu, v[k], 31);
            const int i = base + k * 32 + lane;
            if (i < ny) p[i] = (T)v[k];
        }
    }
}

// cumx_gpu: running sum along x (stride ny).  One CTA per (32-column strip, frame); its WARPS warps split the rows,
// sum their slice, exchange the WARPS x 32 slice totals through shared memory, then rewrite their slice.  32 warps
// give the shortest row walk per warp (small batches: 58 vs 65 us for the whole mask at 20 cm x 8 frames); 16 warps
// let three CTAs share an SM, which wins once there are several waves of strips (10 cm x 32 frames: 0.68 vs 0.73 ms).
template <typename T, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
k_am_scan_x(T *__restrict__ integ, int nx, int ny)
{
    constexpr int kAmScanWarps = WARPS;
    __shared__ int s_part[kAmScanWarps][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    const bool ok = col < ny;
    T *p = integ + (size_t)blockIdx.y * nx * ny + col;
    const int per = (nx + kAmScanWarps - 1) / kAmScanWarps;
    const int r0 = min(nx, warp * per), r1 = min(nx, r0 + per);
    int sum = 0;
    if (ok)
        for (int r = r0; r < r1; ++r) sum += p[(size_t)r * ny];
    s_part[warp][lane] = sum;
    __syncthreads();
    int run = 0;
    for (int w = 0; w < warp; ++w) run += s_part[w][lane];
    if (ok)
        for (int r = r0; r < r1; ++r) {
            run += p[(size_t)r * ny];
            p[(size_t)r * ny] = (T)run;
        }
}

// numpy / numba index rule for the corners: a negative index counts from the end; an index past the end is
// undefined in the reference (no bounds check in numba) and is clamped here.
__device__ __forceinline__ int wrap_index(int i, int n)
{
    if (i < 0) i += n;
    return min(max(i, 0), n - 1);
}

// get_anchors_mask_gpu: mask = (D[maxx,maxy] - D[maxx,miny] - D[minx,maxy] + D[minx,miny]) > 0, the reference's
// four-corner rule as written (it leaves row minx and column miny out of the rectangle).
template <typename T>
__global__ void __launch_bounds__(kAmThreads)
k_am_lookup(const int4 *__restrict__ anchor_coors, long long n_anchors, const T *__restrict__ integ, int n_frames,
            int nx, int ny, uint8_t *__restrict__ mask)
{
    const long long a = (long long)blockIdx.x * kAmThreads + threadIdx.x;
    if (a >= n_anchors) return;
    const int4 c = __ldg(anchor_coors + a);   // (minx, miny, maxx, maxy)
    const int x0 = wrap_index(c.x, nx), y0 = wrap_index(c.y, ny), x1 = wrap_index(c.z, nx), y1 = wrap_index(c.w, ny);
    const int iD = x1 * ny + y1, iA = x0 * ny + y0, iB = x1 * ny + y0, iC = x0 * ny + y1;
    const int f0 = blockIdx.y * kAmFrameGroup, f1 = min(n_frames, f0 + kAmFrameGroup);
    for (int f = f0; f < f1; f += 4) {
        int area[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {          // the sixteen loads of four frames are in flight together
            const T *I = integ + (size_t)min(f + k, f1 - 1) * nx * ny;
            area[k] = (int)I[iD] - (int)I[iB] - (int)I[iC] + (int)I[iA];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (f + k < f1) mask[(size_t)(f + k) * n_anchors + a] = area[k] > 0 ? 1 : 0;
    }
}

template <typename T>
cudaError_t run_anchor_mask(const int32_t *coors, int coors_stride, const int32_t *voxel_num, long long n_rows,
                            int rows_per_frame, int n_frames, int nx, int ny, const int32_t *anchor_coors,
                            long long n_anchors, T *table, uint8_t *mask, cudaStream_t stream)
{
    const size_t cells = (size_t)n_frames * nx * ny;
    cudaError_t e = cudaMemsetAsync(table, 0, (cells * sizeof(T) + 3) & ~(size_t)3, stream);
    if (e != cudaSuccess) return e;
    if (n_rows > 0) {
        const unsigned blocks = (unsigned)((n_rows + kAmThreads - 1) / kAmThreads);
        k_am_count<T><<<blocks, kAmThreads, 0, stream>>>(coors, coors_stride, voxel_num, n_rows, rows_per_frame, n_frames,
                                                         nx, ny, table);
    }
    const long long rows = (long long)n_frames * nx;
    k_am_scan_y<T><<<(unsigned)((rows * 32 + kAmThreads - 1) / kAmThreads), kAmThreads, 0, stream>>>(table, rows, ny);
    const dim3 g_strips((unsigned)((ny + 31) / 32), (unsigned)n_frames);
    if ((long long)g_strips.x * g_strips.y >= 2 * 148)
        k_am_scan_x<T, 16><<<g_strips, 16 * 32, 0, stream>>>(table, nx, ny);
    else
        k_am_scan_x<T, 32><<<g_strips, 32 * 32, 0, stream>>>(table, nx, ny);
    if (n_anchors > 0) {
        const dim3 grid((unsigned)((n_anchors + kAmThreads - 1) / kAmThreads),
                        (unsigned)((n_frames + kAmFrameGroup - 1) / kAmFrameGroup));
        k_am_lookup<T><<<grid, kAmThreads, 0, stream>>>(reinterpret_cast<const int4 *>(anchor_coors), n_anchors, table,
                                                        n_frames, nx, ny, mask);
    }
    return cudaGetLastError();
}

}  // namespace

// `integral` is scratch of n_frames * nx * ny int32; it is used as a uint16 table when no frame can hold more than
// 65535 pillars (pillars_per_frame_bound), which halves the L2 traffic of all four kernels.
cudaError_t launch_anchor_mask(const int32_t *coors, int coors_stride, const int32_t *voxel_num, long long n_rows,
                               int rows_per_frame, int n_frames, int nx, int ny, const int32_t *anchor_coors,
                               long long n_anchors, int32_t *integral, uint8_t *mask, int *table_is_u16,
                               cudaStream_t stream)
{
    const long long bound = rows_per_frame > 0 ? rows_per_frame : n_rows;
    const bool u16 = bound <= 65535;
    if (table_is_u16) *table_is_u16 = u16 ? 1 : 0;
    if (u16)
        return run_anchor_mask<uint16_t>(coors, coors_stride, voxel_num, n_rows, rows_per_frame, n_frames, nx, ny,
                                         anchor_coors, n_anchors, reinterpret_cast<uint16_t *>(integral), mask, stream);
    return run_anchor_mask<int32_t>(coors, coors_stride, voxel_num, n_rows, rows_per_frame, n_frames, nx, ny, anchor_coors,
                                    n_anchors, integral, mask, stream);
}

}  // namespace pf
