// Voxelisation kernels: a data-parallel, bit-exact equivalent of the sequential first-come loop in
// framework/voxel_generator.py:82-106 (reference), batched over independent frames (blockIdx.y = frame).
//
// Order-free definition implemented here (checked on the CPU by oracle/spec_parallel.py):
//   first[c]  = lowest index of an inside point in cell c                          (k_quantize, atomicMin)
//   creator i = first[cell_i] == i ; voxel id = exclusive scan of creator flags     (k_assign, one pass)
//   i*        = the creator whose id equals max_voxels (the loop's `break`), else N
//   slot of i = number of earlier points of the same voxel, from a (tile x voxel) count matrix with a
//               saturating column scan                                           (k_tile_rank, k_column_scan)
//   voxel v keeps the points with i < i* and slot < P, in slot order             (k_slot, k_gather)
//
// Round 2 built and measured two fused forms of K2..K6 (one kernel for ids + ranks + slots over a transposed count
// matrix; one kernel for ids + ranks) -- both correct, both slower than the five launches below:
// profiles/r2_chain_experiments.md.  Kept from them: k_assign numbers its tiles by a per-frame TICKET in CTA start
// order, so the look-back only ever waits on CTAs that are already running (round 1 relied on blockIdx dispatch order);
// every kernel clamps the device-side frame length to the call's max_frame_points.
//
// The file is written against the vocabulary tests/emul/warp_emul.h provides, so that tests/emul/ can compile the very
// same kernel bodies with g++ and hold their logic to the oracle in the build container, which has no GPU
// (PF_HOST_EMUL).  The emulation is test infrastructure; the product only ever runs the nvcc build.
//
// Every kernel is HBM/L2-bound integer work; nothing is a contraction, tensor cores are not involved.
#pragma once

#include "pf_internal.cuh"

namespace pf {

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
template <bool VEC4>
__global__ void __launch_bounds__(256) k_quantize(Params p, Work w, const float *__restrict__ points,
                                                  const int32_t *__restrict__ frame_offsets, int point_stride,
                                                  int lo, int hi)
{
    const int f = blockIdx.y;
    if (lo == 0) pdl_wait();   // first kernel of the chain: everything in front of it on the stream is complete from here on
    pdl_trigger();             // (after the wait, so that no later chain kernel can start before the caller's inputs are visible)
    const int base = frame_offsets[f];
    const int n_all = frame_offsets[f + 1] - base;
    const int n = n_all < hi ? n_all : hi;     // this launch covers [lo, min(n, hi)); hi <= the call's max_frame_points
    const int i0 = lo + blockIdx.x * (256 * kQPer) + threadIdx.x;
    if (lo == 0 && blockIdx.x == 0) {          // per-call state of k_assign: tile tickets, look-back words
        if (threadIdx.x == 0) {
            w.ticket[f] = 0;
            w.istar[f] = 0x7fffffff;   // "no break" default; k_assign lowers it on max_voxels overflow
        }
        for (int t = threadIdx.x; t < p.tiles_cap; t += 256) w.tilecnt[(size_t)f * p.tiles_cap + t] = 0;
    }
    if (i0 >= n) {
        if (lo != 0) pdl_wait();   // every CTA of a dependent grid waits: the grid must not complete before its predecessor
        return;
    }
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
    }
    if (lo != 0) pdl_wait();   // second range: loads and quantisation above ran under the first range's tail; the probes need its grid
#pragma unroll
    for (int k = 0; k < kQPer; ++k)
        seen[k] = c[k] >= 0 ? __ldcg(grid + c[k]) : -1;   // races with other points' atomicMin by design (see above)
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
// Tiles are numbered by a per-frame ticket in CTA start order: the look-back waits only on tiles with a lower
// ticket, i.e. on CTAs that are already running, whatever order the hardware dispatches blockIdx in.
#ifndef PF_ASSIGN_TILE
#define PF_ASSIGN_TILE 4096   // points per k_assign CTA (its own tiling; k_tile_rank keeps kTile).  Measured 2048 /
                              // 4096 / 8192: 19.5 / 16.1 / 18.6 us (20 cm), 29.4 / 22.5 / 22.8 (nuScenes), 82.5 / 72.5 / 72.1 (10 cm)
#endif
constexpr int kATile = PF_ASSIGN_TILE;

__global__ void __launch_bounds__(256) k_assign(Params p, Work w, const int32_t *__restrict__ frame_offsets,
                                                int32_t *__restrict__ voxel_num_out, int max_frame_points)
{
    __shared__ int s_warp[8];
    __shared__ int s_base, s_tile;
    const int f = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_trigger();
    int n = frame_offsets[f + 1] - frame_offsets[f];
    if (n > max_frame_points) n = max_frame_points;   // an over-long frame is truncated, never half-processed
    pdl_wait();                                       // k_quantize: tickets zeroed, grid = first index per cell
    if (tid == 0) s_tile = atomicAdd(w.ticket + f, 1);
    __syncthreads();
    const int t = s_tile;
    const int n_tiles = (n + kATile - 1) / kATile;
    const int start = t * kATile;
    if (n <= 0) {
        if (t == 0 && tid == 0) { w.vnum[f] = 0; if (voxel_num_out) voxel_num_out[f] = 0; }
        return;
    }
    if (start >= n) return;
    const int32_t *cellid = w.cellid + (size_t)f * p.npts_cap;
    int32_t *grid = w.grid + (size_t)f * p.cells;
    int32_t *status = w.tilecnt + (size_t)f * p.tiles_cap;

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
            st_flag(status + t, (int)(((t == 0 ? 2u : 1u) << 30) | (unsigned)block_total));
        int base = 0;
        if (t > 0) {
            int j = t - 1 - lane;   // lane 0 looks at the nearest predecessor
            bool done = false;
            while (!done) {
                unsigned st = 2u << 30;   // "before tile 0": inclusive prefix 0
                if (j >= 0) {
                    // back off between polls: hundreds of warps spin on the lines their predecessors must write
                    while (((st = (unsigned)ld_flag(status + j)) >> 30) == 0u) __nanosleep(50);
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
            if (lane == 0) st_flag(status + t, (int)((2u << 30) | (unsigned)(base + block_total)));
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
                                                               const int32_t *__restrict__ frame_offsets,
                                                               int max_frame_points)
{
    PF_DYN_SHARED(uint8_t, s_tbl);   // [vcols] points seen so far per voxel, this tile
    const int f = blockIdx.y, t = blockIdx.x;
    pdl_trigger();
    int n = frame_offsets[f + 1] - frame_offsets[f];
    if (n > max_frame_points) n = max_frame_points;
    const int start = t * kTile;
    if (start >= n) {
        pdl_wait();   // (every CTA of a dependent grid waits, see k_quantize)
        return;
    }
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the whole table is cleared (not just the n_vox columns in use) so that this can run before k_assign has finished
    for (int j = tid * 16; j < p.vcols; j += kTileWarps * 32 * 16) *reinterpret_cast<uint4 *>(s_tbl + j) = make_uint4(0, 0, 0, 0);
    pdl_wait();
    const int n_vox = w.vnum[f];
    const int istar = w.istar[f];
    const int vbytes = (n_vox + 15) & ~15;

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
// A thread owns 4 adjacent columns (one 32-bit word per row) and walks all rows with 16 independent loads in flight;
// every counter saturates at kCountCap = 255.  (__vaddus4 is emulated on sm_100: 72 instead of 39 us at 10 cm x 32.  Round 2 also tried splitting the rows of
// a column over the warps of a CTA so that every thread issues all its loads at once: same 10 us at 20 cm x 8, 61 instead
// of 38 us at 10 cm x 32 -- profiles/r2_chain_experiments.md E3.)
__global__ void __launch_bounds__(128) k_column_scan(Params p, Work w, const int32_t *__restrict__ frame_offsets,
                                                     int max_frame_points)
{
    const int f = blockIdx.y;
    pdl_trigger();
    int n = frame_offsets[f + 1] - frame_offsets[f];
    if (n > max_frame_points) n = max_frame_points;
    const int n_tiles = (n + kTile - 1) / kTile;
    pdl_wait();
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
                                              int max_frame_points, int reset_grid)
{
    const int f = blockIdx.y;
    pdl_trigger();
    int n = frame_offsets[f + 1] - frame_offsets[f];
    if (n > max_frame_points) n = max_frame_points;
    const int i0 = blockIdx.x * (256 * kQPer) + threadIdx.x;
    pdl_wait();
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
            if (c >= 0) st_same(w.grid + (size_t)f * p.cells + c, kEmpty);   // every point of the cell stores the same word
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

}  // namespace pf
