// TEST INFRASTRUCTURE ONLY.  Runs the voxelisation kernel bodies (csrc/pf_voxel_kernels.cuh: k_quantize, k_assign,
// k_tile_rank, k_column_scan, k_slot) on host threads through warp_emul.h and exposes one batched voxelise call to ctypes, so that
// tests/test_voxel_emul.py can hold the kernel LOGIC (first index per cell, ticketed look-back, cut-off index, in-tile
// ranks, the (tile x voxel) count matrix and its partial rows, slot lists) to the oracle without a GPU.  The final
// per-voxel gather is done here on the host from the kernels' slot lists exactly as k_gather4 reads them.
#define PF_HOST_EMUL 1
#include "pf_voxel_kernels.cuh"

#include <algorithm>

using namespace pf;

namespace {

struct EmulHandle {
    Params p;
    Work w;
    int max_frames;
    std::vector<int32_t> grid, cellid, vr, tilecnt, ticket, slotidx, vcell, vcount, istar, vnum;
    std::vector<uint8_t> cnt;
};

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" {

void *emul_vox_create(const float *offset, const float *voxel_size, const int *grid, int max_voxels, int max_points,
                      int num_features, int max_frames, int max_points_per_frame)
{
    EmulHandle *h = new EmulHandle();
    Params &p = h->p;
    for (int k = 0; k < 3; ++k) p.off[k] = offset[k], p.vs[k] = voxel_size[k];
    p.nx = grid[0], p.ny = grid[1], p.nz = grid[2];
    p.cells = grid[0] * grid[1] * grid[2];
    p.plane = grid[0] * grid[1];
    p.max_voxels = max_voxels, p.P = max_points, p.F = num_features;
    p.npts_cap = (int)align_up((size_t)max_points_per_frame, kTile);
    p.tiles_cap = p.npts_cap / kTile;
    p.vcols = (int)align_up((size_t)max_voxels, 16);
    p.ny_m = (unsigned)((1ull << 32) / (unsigned)p.ny);
    p.nz_m = (unsigned)((1ull << 32) / (unsigned)p.nz);
    p.vbits = 0;
    while ((1 << p.vbits) < max_voxels) ++p.vbits;
    const size_t B = (size_t)max_frames;
    h->max_frames = max_frames;
    h->grid.assign(B * p.cells, kEmpty);
    // everything except the grid starts as garbage on purpose: the kernels must (re)initialise what they read
    h->cellid.assign(B * p.npts_cap, 0x55555555);
    h->vr.assign(B * p.npts_cap, 0x55555555);
    h->tilecnt.assign(B * p.tiles_cap, -1);
    h->ticket.assign(B, 12345);
    h->cnt.assign(B * p.tiles_cap * (size_t)p.vcols, 0xAB);
    h->slotidx.assign(B * max_voxels * (size_t)max_points, -1);
    h->vcell.assign(B * max_voxels, -1);
    h->vcount.assign(B * max_voxels, -1);
    h->vnum.assign(B, -1);
    h->istar.assign(B, -1);
    Work &w = h->w;
    w = Work{};
    w.grid = h->grid.data(), w.cellid = h->cellid.data(), w.vr = h->vr.data(), w.tilecnt = h->tilecnt.data();
    w.ticket = h->ticket.data(), w.cnt = h->cnt.data(), w.slotidx = h->slotidx.data(), w.vcell = h->vcell.data();
    w.vcount = h->vcount.data(), w.vnum = h->vnum.data(), w.istar = h->istar.data();
    return h;
}

void emul_vox_destroy(void *hv) { delete static_cast<EmulHandle *>(hv); }

// One batched voxelise call (slab layout): the launch sequence of launch_voxelize, then the host gather.
// Returns 0, or a negative code when an invariant of the workspace is broken after the call.
int emul_voxelize(void *hv, const float *points, const int *frame_offsets, int n_frames, int max_frame_points,
                  int point_stride, int keep_cell_map, float *voxels, int *coors, int *num_points, int *voxel_num)
{

    EmulHandle *h = static_cast<EmulHandle *>(hv);
    const Params p = h->p;
    const Work w = h->w;
    const int B = n_frames;
    int npts = max_frame_points < 1 ? 1 : max_frame_points;
    const int tiles = (npts + kTile - 1) / kTile;
    constexpr int kPtsPerCta = 256 * kQPer;
    const int split = npts >= 32 * kPtsPerCta ? ((npts / 16 + kPtsPerCta - 1) / kPtsPerCta) * kPtsPerCta : npts;
    for (int pass = 0; pass < 2; ++pass) {
        const int lo = pass == 0 ? 0 : split, hi = pass == 0 ? split : npts;
        if (lo >= hi) continue;
        const dim3 g((unsigned)((hi - lo + kPtsPerCta - 1) / kPtsPerCta), (unsigned)B);
        pf_emul_launch(g, 256, 0, [=] { k_quantize<false>(p, w, points, frame_offsets, point_stride, lo, hi); });
    }
    pf_emul_launch(dim3((unsigned)((npts + kATile - 1) / kATile), (unsigned)B), 256, 0,
                   [=] { k_assign(p, w, frame_offsets, voxel_num, npts); });
    pf_emul_launch(dim3((unsigned)tiles, (unsigned)B), kTileWarps * 32, (size_t)p.vcols,
                   [=] { k_tile_rank(p, w, frame_offsets, npts); });
    pf_emul_launch(dim3((unsigned)((p.vcols / 4 + 127) / 128), (unsigned)B), 128, 0,
                   [=] { k_column_scan(p, w, frame_offsets, npts); });
    pf_emul_launch(dim3((unsigned)((npts + kPtsPerCta - 1) / kPtsPerCta), (unsigned)B), 256, 0,
                   [=] { k_slot(p, w, frame_offsets, npts, keep_cell_map ? 0 : 1); });
    // host gather (what k_gather4 / k_gather do with the slot lists)
    for (int f = 0; f < B; ++f) {
        const int V = w.vnum[f];
        if (V != voxel_num[f]) return -1;
        const int base = frame_offsets[f];
        for (int v = 0; v < V; ++v) {
            const size_t r = (size_t)f * p.max_voxels + v;
            const int cnt = w.vcount[r];
            if (cnt < 1 || cnt > p.P) return -4;
            num_points[r] = cnt;
            const int cell = w.vcell[r];
            const int xy = cell / p.nz;
            coors[3 * r + 0] = xy / p.ny, coors[3 * r + 1] = xy % p.ny, coors[3 * r + 2] = cell % p.nz;
            for (int s = 0; s < p.P; ++s)
                for (int k = 0; k < p.F; ++k)
                    voxels[(r * p.P + s) * p.F + k] =
                        s < cnt ? points[(size_t)(base + w.slotidx[r * p.P + s]) * point_stride + k] : 0.f;
        }
    }
    // invariant the next call relies on: grid all empty unless the map is kept
    if (!keep_cell_map)
        for (int32_t g : h->grid)
            if (g != kEmpty) return -3;
    return 0;
}

// what the canvas fill does with a kept map: every cell back to empty (returns the number of voxel cells seen)
int emul_consume_map(void *hv)
{
    EmulHandle *h = static_cast<EmulHandle *>(hv);
    int n = 0;
    for (int32_t &g : h->grid) {
        if (g < 0) ++n;
        g = kEmpty;
    }
    return n;
}

}  // extern "C"
