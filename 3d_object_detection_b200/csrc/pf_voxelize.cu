// Voxelisation kernels: a data-parallel, bit-exact equivalent of the sequential first-come loop in
// framework/voxel_generator.py:82-106 (reference), batched over independent frames (blockIdx.y = frame).
//
// Order-free definition implemented here (checked on the CPU by oracle/spec_parallel.py):
//   first[c]  = lowest index of an inside point in cell c                       (k_quantize, atomicMin)
//   creator i = first[cell_i] == i ; voxel id = exclusive scan of creator flags  (k_assign, single pass)
//   i*        = the creator whose id equals max_voxels (the loop's `break`), else N
//   slot of i = number of earlier points of the same voxel, from a (tile x voxel) count matrix with a
//               saturating column scan                                           (k_tile_rank, k_column_scan)
//   voxel v keeps the points with i < i* and slot < P, in slot order             (k_slot, k_gather)
//
// Every kernel is HBM/L2-bound integer work; nothing is a contraction, tensor cores are not involved.
#include "pf_internal.cuh"

namespace pf {

// ---------------------------------------------------------------------------------------------------------
__global__ void k_fill_i32(int32_t *__restrict__ dst, int32_t value, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) dst[i] = value;
}

cudaError_t launch_fill_i32(int32_t *dst, int32_t value, size_t n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    size_t blocks = (n + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_fill_i32<<<(unsigned)blocks, threads, 0, stream>>>(dst, value, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------
// K1: quantise + range filter + first index per cell.  A thread owns kQPer points, 256 apart (independent
// load / probe / atomic chains).  float4 loads when the point row is 16 bytes.  Writes cellid[i];
// atomicMin(grid[cell], i).
//
// The L2 serialises atomics per address, a pillar next to the sensor collects thousands of points, and with
// ~1.2 M points in flight they would all hit the same word at once (nuScenes: 48 us for 2.8 M points).  So the
// kernel is launched twice: first over the leading 1/16 of every frame, then over the rest, and every point
// probes the cell first (an L2 read; a cell only ever decreases, so a stale value can only cause a redundant
// atomic): in the second launch any busy cell already holds an index from the first range and is left alone.
constexpr int kQPer = 4;

template <bool VEC4>
__global__ void __launch_bounds__(256) k_quantize(Params p, Work w, const float *__restrict__ points,
                                                  const int32_t *__restrict__ frame_offsets, int point_stride,
                                                  int lo, int hi)
{
    const int f = blockIdx.y;
    const int base = frame_offsets[f];
    const int n_all = frame_offsets[f + 1] - base;
    const int n = n_all < hi ? n_all : hi;     // this launch covers [lo, min(n, hi))
    const int i0 = lo + blockIdx.x * (256 * kQPer) + threadIdx.x;
    if (lo == 0 && blockIdx.x == 0) {
        if (threadIdx.x == 0) w.istar[f] = n_all;   // "no break" default; k_assign lowers it on max_voxels overflow
        for (int t = threadIdx.x; t < p.tiles_cap; t += 256)
            w.tilecnt[(size_t)f * p.tiles_cap + t] = 0;   // look-back status words of k_assign: "not ready"
    }
    if (i0 >= n) return;
    int c[kQPer];
    float x[kQPer], y[kQPer], z[kQPer];
#pragma unroll
    for (int k = 0; k < kQPer; ++k) {
        const int i = i0 + k * 256;
        x[k] = y[k] = z[k] = __int_as_float(0x7fc00000);   // NaN: outside
        if (i < n) {
            if (VEC4) {
                const float4 q = __ldg(reinterpret_cast<const float4 *>(points) + (size_t)(base + i));
                x[k] = q.x, y[k] = q.y, z[k] = q.z;
            } else {
                const float *q = points + (size_t)(base + i) * point_stride;
                x[k] = __ldg(q), y[k] = __ldg(q + 1), z[k] = __ldg(q + 2);
            }
        }
    }
    int32_t *grid = w.grid + (size_t)f * p.cells;
    int seen[kQPer];
#pragma unroll
    for (int k = 0; k < kQPer; ++k) {
        const int i = i0 + k * 256;
        c[k] = cell_of_point(x[k], y[k], z[k], p);
        if (i < n) w.cellid[(size_t)f * p.npts_cap + i] = c[k];
        seen[k] = c[k] >= 0 ? __ldcg(grid + c[k]) : -1;
    }
#pragma unroll
    for (int k = 0; k < kQPer; ++k) {
        const int i = i0 + k * 256;
        if (seen[k] > i) atomicMin(grid + c[k], i);
    }
}

// ---------------------------------------------------------------------------------------------------------
// K2: voxel ids in ONE pass over the points.  A point is the creator of its cell's voxel iff it holds the
// cell's first index; the voxel id is the exclusive scan of the creator flags in point order.  Tiles of kTile
// points scan locally (ballot/shuffle) and chain their totals with a decoupled look-back over the per-tile
// status words (tilecnt: flag << 30 | count; 1 = tile total, 2 = inclusive prefix; zeroed by k_quantize).
// Creators below max_voxels record their cell and overwrite grid[cell] with ~id; the creator whose id equals
// max_voxels publishes the cut-off index i* (the reference's `break`).
// Look-back waits only on tiles with a lower blockIdx.x of the same frame, which are dispatched earlier.
__device__ __forceinline__ unsigned ld_status(const int32_t *p)
{
    return *reinterpret_cast<const volatile unsigned *>(p);
}

#ifndef PF_ASSIGN_TILE
#define PF_ASSIGN_TILE 4096   // points per k_assign CTA (its own tiling; k_tile_rank keeps kTile).  Measured 2048 /
                              // 4096 / 8192: 19.5 / 16.1 / 18.6 us (20 cm), 29.4 / 22.5 / 22.8 (nuScenes), 82.5 / 72.5 / 72.1 (10 cm)
#endif
constexpr int kATile = PF_ASSIGN_TILE;

__global__ void __launch_bounds__(256) k_assign(Params p, Work w, const int32_t *__restrict__ frame_offsets,
                                                int32_t *__restrict__ voxel_num_out)
{
    const int f = blockIdx.y, t = blockIdx.x;
    const int n = frame_offsets[f + 1] - frame_offsets[f];
    const int n_tiles = (n + kATile - 1) / kATile;
    const int start = t * kATile;
    if (n == 0) {
        if (t == 0 && threadIdx.x == 0) { w.vnum[f] = 0; if (voxel_num_out) voxel_num_out[f] = 0; }
        return;
    }
    if (start >= n) return;
    const int32_t *cellid = w.cellid + (size_t)f * p.npts_cap;
    int32_t *grid = w.grid + (size_t)f * p.cells;
    int32_t *status = w.tilecnt + (size_t)f * p.tiles_cap;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    __shared__ int s_warp[8];
    __shared__ int s_base;

    // thread owns PER consecutive points (the per-point scratch is padded to a multiple of kTile, not of kATile)
    constexpr int PER = kATile / 256;
    const int i0 = start + tid * PER;
    int cells[PER];
#pragma unroll
    for (int j = 0; j < PER; j += 4) {
        int4 a = make_int4(-1, -1, -1, -1);
        if (i0 + j < p.npts_cap) a = *reinterpret_cast<const int4 *>(cellid + i0 + j);
        cells[j] = a.x, cells[j + 1] = a.y, cells[j + 2] = a.z, cells[j + 3] = a.w;
    }
    int first[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k) first[k] = (i0 + k < n && cells[k] >= 0) ? __ldcg(grid + cells[k]) : -1;
    unsigned flags = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k)
        if (first[k] == i0 + k) flags |= 1u << k;
    const int mine = __popc(flags);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int warp_off = 0, block_total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int sv = s_warp[k];
        if (k < warp) warp_off += sv;
        block_total += sv;
    }
    if (warp == 0) {
        // publish this tile's total first so that no successor ever waits on our own look-back
        if (lane == 0)
            *reinterpret_cast<volatile unsigned *>(status + t) = ((t == 0 ? 2u : 1u) << 30) | (unsigned)block_total;
        int base = 0;
        if (t > 0) {
            int j = t - 1 - lane;   // lane 0 looks at the nearest predecessor
            bool done = false;
            while (!done) {
                unsigned st = 2u << 30;   // "before tile 0": inclusive prefix 0
                if (j >= 0) {
                    // back off between polls: hundreds of warps spin on the lines their predecessors must write
                    while (((st = ld_status(status + j)) >> 30) == 0u) __nanosleep(50);
                }
                const unsigned inc_mask = __ballot_sync(0xffffffffu, (st >> 30) == 2u);
                int val = (int)(st & 0x3fffffffu);
                if (inc_mask) {
                    const int nearest = __ffs(inc_mask) - 1;
                    if (lane > nearest) val = 0;
                    done = true;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
                base += val;
                j -= 32;
            }
            if (lane == 0) *reinterpret_cast<volatile unsigned *>(status + t) = (2u << 30) | (unsigned)(base + block_total);
        }
        if (lane == 0) s_base = base;
    }
    __syncthreads();
    const int tile_base = s_base;

    int rank = tile_base + warp_off + incl - mine;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        if (flags & (1u << k)) {
            if (rank < p.max_voxels) {
                w.vcell[(size_t)f * p.max_voxels + rank] = cells[k];
                grid[cells[k]] = ~rank;
            } else if (rank == p.max_voxels) {
                w.istar[f] = i0 + k;
            }
            ++rank;
        }
    }
    if (t == n_tiles - 1 && tid == 0) {
        const int total = tile_base + block_total;
        const int v = total < p.max_voxels ? total : p.max_voxels;
        w.vnum[f] = v;
        if (voxel_num_out) voxel_num_out[f] = v;
    }
}

// ---------------------------------------------------------------------------------------------------------
// K4: in-tile ranks.  A tile is kTile consecutive points; warp w owns points [w*256, (w+1)*256) of the tile
// and walks them in kTileSteps steps of 32, so (warp, step, lane) order == point order.  Lanes of one step
// that fall in the same voxel are grouped (ballots over the id bits); the group's lowest lane bumps the voxel's byte
// counter in shared memory and the group derives its ranks from the old value.  Warps take turns (warp 0
// first) so counters are bumped in point order without atomics.  Result: vr[i] = (voxel << 8 | rank in
// tile) and one row of the (tile x voxel) count matrix, written densely so the matrix never needs clearing.
__global__ void __launch_bounds__(kTileWarps * 32) k_tile_rank(Params p, Work w,
                                                               const int32_t *__restrict__ frame_offsets)
{
    extern __shared__ __align__(16) uint8_t s_tbl[];   // [vcols] points seen so far per voxel, this tile
    const int f = blockIdx.y, t = blockIdx.x;
    const int n = frame_offsets[f + 1] - frame_offsets[f];
    const int start = t * kTile;
    if (start >= n) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_vox = w.vnum[f];
    const int istar = w.istar[f];
    const int vbytes = (n_vox + 15) & ~15;
    for (int j = tid * 16; j < vbytes; j += kTileWarps * 32 * 16) *reinterpret_cast<uint4 *>(s_tbl + j) = make_uint4(0, 0, 0, 0);

    const int32_t *cellid = w.cellid + (size_t)f * p.npts_cap;
    const int32_t *grid = w.grid + (size_t)f * p.cells;
    int vid[kTileSteps];
    unsigned grp[kTileSteps];
    const int i_first = start + warp * (32 * kTileSteps) + lane;
#pragma unroll
    for (int s = 0; s < kTileSteps; ++s) {
        const int i = i_first + s * 32;
        int c = -1;
        if (i < n && i < istar) c = cellid[i];
        vid[s] = c;
    }
#pragma unroll
    for (int s = 0; s < kTileSteps; ++s) {
        const int c = vid[s];
        vid[s] = (c >= 0) ? ~grid[c] : -1;   // every inside point below i* sits in a cell that owns a voxel
    }
    // Lanes of a step with the same voxel: one ballot per bit of the voxel id, each lane keeps the lanes that agree
    // with it on every bit.  (match.any does this in one instruction, but its cost grows with the number of distinct
    // keys -- ~330 cycles for 32 different voxels -- and it was half of this kernel's time; vbits ballots are ~60.)
#pragma unroll
    for (int s = 0; s < kTileSteps; ++s) {
        const bool ok = vid[s] >= 0;
        unsigned m = __ballot_sync(0xffffffffu, ok);
        for (int b = 0; b < p.vbits; ++b) {
            const bool one = (vid[s] >> b) & 1;
            const unsigned bal = __ballot_sync(0xffffffffu, one);
            m &= one ? bal : ~bal;
        }
        grp[s] = ok ? m : (1u << lane);
    }
    __syncthreads();   // table zeroed

    int rnk[kTileSteps];
    for (int turn = 0; turn < kTileWarps; ++turn) {
        if (warp == turn) {
#pragma unroll
            for (int s = 0; s < kTileSteps; ++s) {
                const int leader = __ffs(grp[s]) - 1;
                int base = 0;
                if (vid[s] >= 0 && lane == leader) {
                    base = s_tbl[vid[s]];
                    const int nxt = base + __popc(grp[s]);
                    s_tbl[vid[s]] = (uint8_t)(nxt < kCountCap ? nxt : kCountCap);
                }
                base = __shfl_sync(0xffffffffu, base, leader);
                const int r = base + __popc(grp[s] & lanemask_lt());
                rnk[s] = r < kCountCap ? r : kCountCap;
                __syncwarp();
            }
        }
        __syncthreads();
    }

    int32_t *vr = w.vr + (size_t)f * p.npts_cap;
#pragma unroll
    for (int s = 0; s < kTileSteps; ++s) {
        const int i = i_first + s * 32;
        if (i < n) vr[i] = vid[s] >= 0 ? ((vid[s] << 8) | rnk[s]) : -1;
    }
    uint8_t *row = w.cnt + ((size_t)f * p.tiles_cap + t) * p.vcols;
    for (int j = tid * 16; j < vbytes; j += kTileWarps * 32 * 16)
        *reinterpret_cast<uint4 *>(row + j) = *reinterpret_cast<const uint4 *>(s_tbl + j);
}

// ---------------------------------------------------------------------------------------------------------
// K5: saturating exclusive scan down every voxel column of the count matrix (in place), totals -> vcount.
// A thread owns 4 adjacent columns (one 32-bit word per row); rows are independent loads.
__global__ void __launch_bounds__(128) k_column_scan(Params p, Work w, const int32_t *__restrict__ frame_offsets)
{
    const int f = blockIdx.y;
    const int n = frame_offsets[f + 1] - frame_offsets[f];
    const int n_tiles = (n + kTile - 1) / kTile;
    const int n_vox = w.vnum[f];
    const int col = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (col >= n_vox) return;
    uint32_t *m = reinterpret_cast<uint32_t *>(w.cnt + (size_t)f * p.tiles_cap * p.vcols + col);
    const size_t row_words = p.vcols / 4;
    int run[4] = {0, 0, 0, 0};
    constexpr int U = 16;
    for (int t0 = 0; t0 < n_tiles; t0 += U) {
        uint32_t v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = (t0 + u < n_tiles) ? m[(size_t)(t0 + u) * row_words] : 0u;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (t0 + u < n_tiles) {
                uint32_t out = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    out |= (uint32_t)run[k] << (8 * k);
                    const int s = run[k] + (int)((v[u] >> (8 * k)) & 0xffu);
                    run[k] = s < kCountCap ? s : kCountCap;
                }
                m[(size_t)(t0 + u) * row_words] = out;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (col + k < n_vox) w.vcount[(size_t)f * p.max_voxels + col + k] = run[k] < p.P ? run[k] : p.P;
}

// ---------------------------------------------------------------------------------------------------------
// K6: slot of every surviving point = scanned count of its (tile, voxel) + its in-tile rank; the first P
// slots of a voxel record the point index.  Also puts the touched grid cells back to kEmpty so the grid
// never needs a memset (skipped when the cell map is kept for pf_scatter_from_map, which resets it).
__global__ void __launch_bounds__(256) k_slot(Params p, Work w, const int32_t *__restrict__ frame_offsets,
                                              int reset_grid)
{
    const int f = blockIdx.y;
    const int n = frame_offsets[f + 1] - frame_offsets[f];
    const int i0 = blockIdx.x * (256 * kQPer) + threadIdx.x;
    if (i0 >= n) return;
    int vr[kQPer], cnt[kQPer];
#pragma unroll
    for (int k = 0; k < kQPer; ++k) {
        const int i = i0 + k * 256;
        vr[k] = i < n ? w.vr[(size_t)f * p.npts_cap + i] : -1;
    }
    if (reset_grid) {
#pragma unroll
        for (int k = 0; k < kQPer; ++k) {
            const int i = i0 + k * 256;
            const int c = i < n ? w.cellid[(size_t)f * p.npts_cap + i] : -1;
            if (c >= 0) w.grid[(size_t)f * p.cells + c] = kEmpty;
        }
    }
#pragma unroll
    for (int k = 0; k < kQPer; ++k) {
        const int i = i0 + k * 256;
        cnt[k] = vr[k] >= 0 ? (int)w.cnt[((size_t)f * p.tiles_cap + i / kTile) * p.vcols + (vr[k] >> 8)] : 0;
    }
#pragma unroll
    for (int k = 0; k < kQPer; ++k) {
        if (vr[k] < 0) continue;
        const int v = vr[k] >> 8, slot = cnt[k] + (vr[k] & 0xff);
        if (slot < p.P) w.slotidx[((size_t)f * p.max_voxels + v) * p.P + slot] = i0 + k * 256;
    }
}

// ---------------------------------------------------------------------------------------------------------
// K7: per-voxel gather + pillar decoration (networks/pointpillars8_shared.py:32-54 of the reference).
// GROUP lanes (16 when P <= 16, else 32) own one voxel; a 256-thread CTA stages 256/GROUP voxels in shared
// memory and then streams the voxel rows and the decorated rows out with 16-byte stores, so every output
// byte -- padding zeros included -- is written exactly once and fully coalesced.
//   FROM_POINTS = true : rows come from the point cloud through slotidx (fused voxelise path)
//   FROM_POINTS = false: rows come from an existing voxels tensor (stand-alone pf_decorate)

__device__ __forceinline__ void copy_out(float *__restrict__ gdst, const float *ssrc, int count, int phase)
{
    // gdst and ssrc share the same 16-byte phase (ssrc = aligned base + phase floats)
    int head = (4 - phase) & 3;
    if (head > count) head = count;
    if ((int)threadIdx.x < head) gdst[threadIdx.x] = ssrc[threadIdx.x];
    const int nvec = (count - head) >> 2;
    float4 *gd = reinterpret_cast<float4 *>(gdst + head);
    const float4 *sd = reinterpret_cast<const float4 *>(ssrc + head);
    for (int j = threadIdx.x; j < nvec; j += blockDim.x) gd[j] = sd[j];
    const int done = head + (nvec << 2);
    if ((int)threadIdx.x < count - done) gdst[done + threadIdx.x] = ssrc[done + threadIdx.x];
}

template <int GROUP, bool FROM_POINTS, bool VEC4>
__global__ void __launch_bounds__(256) k_gather(Params p, Work w, GatherArgs a)
{
    constexpr int VPB = 256 / GROUP;
    extern __shared__ __align__(16) float s_mem[];
    const int P = p.P, F = p.F, FO = p.F + 5;
    float *s_vox = s_mem;                                 // [4 + VPB*P*F]
    float *s_feat = s_mem + ((VPB * P * F + 4 + 3) & ~3); // [4 + VPB*P*FO]

    const int f = FROM_POINTS ? blockIdx.y : 0;
    const long long n_vox = FROM_POINTS ? (long long)w.vnum[f] : a.n_voxels;
    const long long v0 = (long long)blockIdx.x * VPB;
    if (v0 >= n_vox) return;
    const int nv = (int)((n_vox - v0) < VPB ? (n_vox - v0) : VPB);

    long long row0 = v0;   // first output row of this CTA
    int pt_base = 0;
    if (FROM_POINTS) {
        pt_base = a.frame_offsets[f];
        if (a.compact) {
            long long s = 0;
            for (int k = 0; k < f; ++k) s += w.vnum[k];
            row0 += s;
        } else {
            row0 += (long long)f * p.max_voxels;
        }
    }
    const int phase_v = (int)((row0 * P * F) & 3);
    const int phase_f = (int)((row0 * P * FO) & 3);

    const int tid = threadIdx.x;
    const int g = tid / GROUP, gl = tid % GROUP;
    const bool active = g < nv;
    const long long v = v0 + g;

    int cnt = 0, cx = 0, cy = 0, cz = 0;
    if (active) {
        if (FROM_POINTS) {
            cnt = w.vcount[(size_t)f * p.max_voxels + v];
            const int cell = w.vcell[(size_t)f * p.max_voxels + v];
            cz = cell % p.nz;
            const int xy = cell / p.nz;
            cy = xy % p.ny;
            cx = xy / p.ny;
        } else {
            cnt = a.num_in[v];
            cx = a.coors_in[v * a.coors_stride + 0];
            cy = a.coors_in[v * a.coors_stride + 1];
        }
    }
    // stage the voxel rows, accumulate the xyz sums
    float sx = 0.f, sy = 0.f, sz = 0.f;
    float *vrow = s_vox + phase_v + (size_t)g * P * F;
    if (active) {
        for (int r = gl; r < P; r += GROUP) {
            float *dst = vrow + r * F;
            if (FROM_POINTS) {
                if (r < cnt) {
                    const int idx = w.slotidx[((size_t)f * p.max_voxels + v) * P + r];
                    if (VEC4) {
                        const float4 q = __ldg(reinterpret_cast<const float4 *>(a.points) + (size_t)(pt_base + idx));
                        dst[0] = q.x, dst[1] = q.y, dst[2] = q.z, dst[3] = q.w;
                        sx += q.x, sy += q.y, sz += q.z;
                    } else {
                        const float *q = a.points + (size_t)(pt_base + idx) * a.point_stride;
                        for (int k = 0; k < F; ++k) dst[k] = __ldg(q + k);
                        sx += dst[0], sy += dst[1], sz += dst[2];
                    }
                } else {
                    for (int k = 0; k < F; ++k) dst[k] = 0.f;
                }
            } else {
                const float *q = a.voxels_in + ((size_t)v * P + r) * F;
                for (int k = 0; k < F; ++k) dst[k] = __ldg(q + k);
                sx += dst[0], sy += dst[1], sz += dst[2];   // the reference sums every slot (:32)
            }
        }
    }
#pragma unroll
    for (int o = GROUP / 2; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o);
    }
    if (active && a.features) {
        const float fc = (float)cnt;
        const float mx = __fdiv_rn(sx, fc), my = __fdiv_rn(sy, fc), mz = __fdiv_rn(sz, fc);
        // pillar centre: float(c) * v rounded, then + offset rounded -- two fp32 ops, no FMA
        const float ccx = __fadd_rn(__fmul_rn((float)cx, a.vx), a.x_off);
        const float ccy = __fadd_rn(__fmul_rn((float)cy, a.vy), a.y_off);
        float *frow = s_feat + phase_f + (size_t)g * P * FO;
        for (int r = gl; r < P; r += GROUP) {
            const float *src = vrow + r * F;
            float *dst = frow + r * FO;
            if (r < cnt) {
                for (int k = 0; k < F; ++k) dst[k] = src[k];
                dst[F + 0] = __fsub_rn(src[0], mx);
                dst[F + 1] = __fsub_rn(src[1], my);
                dst[F + 2] = __fsub_rn(src[2], mz);
                dst[F + 3] = __fsub_rn(src[0], ccx);
                dst[F + 4] = __fsub_rn(src[1], ccy);
            } else {
                for (int k = 0; k < FO; ++k) dst[k] = 0.f;
            }
        }
    }
    if (FROM_POINTS && active && gl == 0) {
        const long long row = row0 + g;
        if (a.coors) {
            int32_t *c = a.coors + row * a.coors_stride;
            c[0] = cx, c[1] = cy, c[2] = cz;
            if (a.coors_stride >= 4) c[3] = f;
        }
        if (a.num_points) a.num_points[row] = cnt;
    }
    __syncthreads();
    if (FROM_POINTS && a.voxels) copy_out(a.voxels + (size_t)row0 * P * F, s_vox + phase_v, nv * P * F, phase_v);
    if (a.features) copy_out(a.features + (size_t)row0 * P * FO, s_feat + phase_f, nv * P * FO, phase_f);
}

// ---------------------------------------------------------------------------------------------------------
// host-side launchers
static size_t gather_smem_bytes(int group, int P, int F)
{
    const int vpb = 256 / group;
    const size_t a = ((size_t)vpb * P * F + 4 + 3) & ~(size_t)3;
    const size_t b = (size_t)vpb * P * (F + 5) + 4;
    return (a + b) * sizeof(float);
}

template <int GROUP, bool FROM_POINTS, bool VEC4>
static cudaError_t run_gather(const Params &p, const Work &w, const GatherArgs &a, dim3 grid, cudaStream_t stream)
{
    const size_t smem = gather_smem_bytes(GROUP, p.P, p.F);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_gather<GROUP, FROM_POINTS, VEC4>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_gather<GROUP, FROM_POINTS, VEC4><<<grid, 256, smem, stream>>>(p, w, a);
    return cudaGetLastError();
}

template <bool FROM_POINTS>
static cudaError_t dispatch_gather(const Params &p, const Work &w, const GatherArgs &a, dim3 grid_g16, dim3 grid_g32,
                                   bool vec4, cudaStream_t stream)
{
    if (p.P <= 16) {
        return vec4 ? run_gather<16, FROM_POINTS, true>(p, w, a, grid_g16, stream)
                    : run_gather<16, FROM_POINTS, false>(p, w, a, grid_g16, stream);
    }
    return vec4 ? run_gather<32, FROM_POINTS, true>(p, w, a, grid_g32, stream)
                : run_gather<32, FROM_POINTS, false>(p, w, a, grid_g32, stream);
}

cudaError_t launch_voxelize(const Params &p, const Work &w, const VoxelizeArgs &a, cudaStream_t stream)
{
    const int B = a.n_frames;
    if (B <= 0) return cudaSuccess;
    int npts = a.max_frame_points;
    if (npts < 1) npts = 1;
    const int tiles = (npts + kTile - 1) / kTile;
    const bool vec4 = (a.point_stride == 4) && (p.F == 4) && ((reinterpret_cast<uintptr_t>(a.points) & 15) == 0);
    constexpr int kPtsPerCta = 256 * kQPer;
    const dim3 g_pts((npts + kPtsPerCta - 1) / kPtsPerCta, B), g_tiles(tiles, B);
    // leading 1/16 of every frame first (see k_quantize); small frames go in one launch
    const int split = npts >= 32 * kPtsPerCta ? ((npts / 16 + kPtsPerCta - 1) / kPtsPerCta) * kPtsPerCta : npts;
    mark(w, kEvStart, stream);
    for (int pass = 0; pass < 2; ++pass) {
        const int lo = pass == 0 ? 0 : split, hi = pass == 0 ? split : npts;
        if (lo >= hi) continue;
        const dim3 g((hi - lo + kPtsPerCta - 1) / kPtsPerCta, B);
        if (vec4) k_quantize<true><<<g, 256, 0, stream>>>(p, w, a.points, a.frame_offsets, a.point_stride, lo, hi);
        else      k_quantize<false><<<g, 256, 0, stream>>>(p, w, a.points, a.frame_offsets, a.point_stride, lo, hi);
    }
    mark(w, kEvQuantize, stream);
    k_assign<<<dim3((npts + kATile - 1) / kATile, B), 256, 0, stream>>>(p, w, a.frame_offsets, a.voxel_num);
    mark(w, kEvAssign, stream);
    if ((size_t)p.vcols > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_tile_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, p.vcols);
        if (e != cudaSuccess) return e;
    }
    k_tile_rank<<<g_tiles, kTileWarps * 32, p.vcols, stream>>>(p, w, a.frame_offsets);
    mark(w, kEvTileRank, stream);
    k_column_scan<<<dim3((p.vcols / 4 + 127) / 128, B), 128, 0, stream>>>(p, w, a.frame_offsets);
    mark(w, kEvColumnScan, stream);
    k_slot<<<g_pts, 256, 0, stream>>>(p, w, a.frame_offsets, a.keep_cell_map ? 0 : 1);
    mark(w, kEvSlot, stream);

    GatherArgs ga{};
    ga.points = a.points, ga.frame_offsets = a.frame_offsets, ga.point_stride = a.point_stride;
    ga.compact = a.compact, ga.coors_stride = a.coors_stride;
    ga.vx = a.vx, ga.vy = a.vy, ga.x_off = a.x_off, ga.y_off = a.y_off;
    ga.voxels = a.voxels, ga.coors = a.coors, ga.num_points = a.num_points, ga.features = a.features;
    ga.pfn_w = a.pfn_w, ga.pfn_scale = a.pfn_scale, ga.pfn_shift = a.pfn_shift, ga.pfn_out = a.pfn_out;
    cudaError_t e = vec4 ? launch_gather4(p, w, ga, B, true, stream) : cudaErrorNotSupported;
    if (e == cudaErrorNotSupported && a.pfn_w) return e;   // the fused PFN exists only on the F = 4 fast path
    if (e == cudaErrorNotSupported)   // generic F / stride / alignment
        e = dispatch_gather<true>(p, w, ga, dim3((p.max_voxels + 15) / 16, B), dim3((p.max_voxels + 7) / 8, B), vec4,
                                  stream);
    if (e != cudaSuccess) return e;
    mark(w, kEvGather, stream);
    return cudaGetLastError();
}

cudaError_t launch_decorate(const float *voxels, const int32_t *num, const int32_t *coors, int coors_stride,
                            int64_t n_voxels, int P, int F, float vx, float vy, float x_off, float y_off,
                            float *features, cudaStream_t stream, const float *pfn_w, const float *pfn_scale,
                            const float *pfn_shift, float *pfn_out)
{
    if (n_voxels <= 0) return cudaSuccess;
    Params p{};
    p.P = P, p.F = F, p.nx = p.ny = p.nz = 1;
    Work w{};
    GatherArgs ga{};
    ga.voxels_in = voxels, ga.num_in = num, ga.coors_in = coors, ga.coors_stride = coors_stride;
    ga.n_voxels = n_voxels;
    ga.vx = vx, ga.vy = vy, ga.x_off = x_off, ga.y_off = y_off;
    ga.features = features;
    ga.pfn_w = pfn_w, ga.pfn_scale = pfn_scale, ga.pfn_shift = pfn_shift, ga.pfn_out = pfn_out;
    const cudaError_t e4 = launch_gather4(p, w, ga, 1, false, stream);
    if (e4 != cudaErrorNotSupported || pfn_w) return e4;   // the fused PFN exists only on the F = 4 fast path
    const unsigned b16 = (unsigned)((n_voxels + 15) / 16), b32 = (unsigned)((n_voxels + 7) / 8);
    return dispatch_gather<false>(p, w, ga, dim3(b16, 1), dim3(b32, 1), false, stream);
}

}  // namespace pf
