"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Data-parallel restatement of the sequential voxeliser.

The numba loop (framework/voxel_generator.py:82-106) is first-come/first-served and stops at the first
new cell after ``max_voxels`` cells exist.  The CUDA path cannot walk points in order, so it implements
the equivalent order-free definition below.  This module states that definition twice in
numpy:

* :func:`voxelize_sorted` -- the specification (stable sort by cell).
* :func:`voxelize_tiled`  -- the *algorithm the kernels run*: per-cell first index by min-reduction,
  creator flags + exclusive scan for voxel ids, cut-off index ``i*``, and in-voxel slots from a
  (tile x voxel) count matrix with a saturating column scan.  It exists so the kernel algorithm can be
  checked on a CPU against the C oracle and the reference's golden vectors before any GPU is involved.

Both must be bit-identical to oracle.generate(); tests/test_oracle.py checks it.
"""
from __future__ import annotations

import numpy as np

EMPTY = np.iinfo(np.int32).max


def quantize(geom, points):
    """Linear cell id per point, -1 if outside (F4, F7): fp32 sub, fp32 div, floor, range test."""
    pts = np.asarray(points, np.float32)
    n = pts.shape[0]
    gs = geom.grid_size.astype(np.int64)
    with np.errstate(invalid="ignore", over="ignore", divide="ignore"):
        q = np.floor((pts[:, :3] - geom.offset.astype(np.float32)) / geom.voxel_size.astype(np.float32))
        inside = np.all((q >= 0) & (q < gs.astype(np.float32)), axis=1)   # NaN/Inf compare false
    c = np.where(inside[:, None], q, 0).astype(np.int64)
    cell = (c[:, 0] * gs[1] + c[:, 1]) * gs[2] + c[:, 2]
    return np.where(inside, cell, -1), n


def _coords_of(cell, geom):
    gs = geom.grid_size.astype(np.int64)
    z = cell % gs[2]
    y = (cell // gs[2]) % gs[1]
    x = cell // (gs[2] * gs[1])
    return np.stack([x, y, z], axis=1).astype(np.int32)


def voxelize_sorted(geom, points, max_voxels, max_points):
    pts = np.ascontiguousarray(points, np.float32)
    cell, n = quantize(geom, pts)
    f = pts.shape[1]
    idx = np.nonzero(cell >= 0)[0]
    if idx.size == 0:
        return (np.zeros((0, max_points, f), np.float32), np.zeros((0, 3), np.int32), np.zeros((0,), np.int32))
    uniq, first_pos = np.unique(cell[idx], return_index=True)
    first = idx[first_pos]                                # first point index of every touched cell
    order = np.argsort(first, kind="stable")              # voxel id = rank of first index
    cells_by_rank = uniq[order]
    first_by_rank = first[order]
    v = min(len(order), max_voxels)
    istar = first_by_rank[max_voxels] if len(order) > max_voxels else n
    rank_of_cell = {int(c): r for r, c in enumerate(cells_by_rank[:v])}
    voxels = np.zeros((v, max_points, f), np.float32)
    num = np.zeros((v,), np.int32)
    for i in idx[idx < istar]:
        r = rank_of_cell[int(cell[i])]
        if num[r] < max_points:
            voxels[r, num[r]] = pts[i]
            num[r] += 1
    return voxels, _coords_of(cells_by_rank[:v], geom), num


def voxelize_tiled(geom, points, max_voxels, max_points, tile=2048, count_cap=255):
    """The kernel pipeline, stage by stage (names match 3d_object_detection_b200/csrc/pf_voxelize.cu)."""
    pts = np.ascontiguousarray(points, np.float32)
    f = pts.shape[1]
    cell, n = quantize(geom, pts)                                     # k_quantize: cell ids ...
    first = np.full(geom.cells, EMPTY, np.int64)
    ins = np.nonzero(cell >= 0)[0]
    np.minimum.at(first, cell[ins], ins)                              # ... and atomicMin(first[cell], i)
    creator = np.zeros(n, bool)
    creator[ins] = first[cell[ins]] == ins                            # k_creator_count / k_assign
    n_tiles = (n + tile - 1) // tile
    tile_cnt = np.array([creator[t * tile:(t + 1) * tile].sum() for t in range(n_tiles)], np.int64)
    tile_base = np.concatenate([[0], np.cumsum(tile_cnt)])[:n_tiles] if n_tiles else np.zeros(0, np.int64)
    rank = np.full(n, -1, np.int64)
    for t in range(n_tiles):
        sl = slice(t * tile, min(n, (t + 1) * tile))
        rank[sl] = np.where(creator[sl], tile_base[t] + np.cumsum(creator[sl]) - creator[sl], -1)
    total = int(creator.sum())
    v = min(total, max_voxels)
    over = np.nonzero(rank == max_voxels)[0]
    istar = int(over[0]) if over.size else n
    grid = first.copy()
    keepers = np.nonzero(creator & (rank < max_voxels))[0]
    grid[cell[keepers]] = -1 - rank[keepers]                          # encoded voxel id (negative)
    coors = np.zeros((v, 3), np.int32)
    coors[rank[keepers]] = _coords_of(cell[keepers], geom)
    # k_tile_rank: voxel id + in-tile rank per point, (tile x voxel) count matrix, saturating
    valid = (cell >= 0) & (np.arange(n) < istar)
    vid = np.full(n, -1, np.int64)
    vid[valid] = -1 - grid[cell[valid]]
    assert np.all(vid[valid] >= 0) and np.all(vid[valid] < v)
    cnt = np.zeros((n_tiles, max(v, 1)), np.int64)
    r_in_tile = np.zeros(n, np.int64)
    for t in range(n_tiles):
        run = {}
        for i in range(t * tile, min(n, (t + 1) * tile)):
            if valid[i]:
                k = run.get(int(vid[i]), 0)
                r_in_tile[i] = min(k, count_cap)
                run[int(vid[i])] = min(k + 1, count_cap)
        for key, k in run.items():
            cnt[t, key] = k
    # k_column_scan: exclusive saturating prefix over tiles; totals -> num
    pre = np.zeros_like(cnt)
    run = np.zeros(cnt.shape[1], np.int64)
    for t in range(n_tiles):
        pre[t] = run
        run = np.minimum(run + cnt[t], count_cap)
    num = np.minimum(run[:v], max_points).astype(np.int32)
    # k_slot: slot = pre[tile][voxel] + in-tile rank, keep if < P; k_gather copies points
    slot_idx = np.full((v, max_points), -1, np.int64)
    for i in np.nonzero(valid)[0]:
        s = pre[i // tile, vid[i]] + r_in_tile[i]
        if s < max_points:
            slot_idx[vid[i], s] = i
    voxels = np.zeros((v, max_points, f), np.float32)
    for vv in range(v):
        k = num[vv]
        assert np.all(slot_idx[vv, :k] >= 0)
        voxels[vv, :k] = pts[slot_idx[vv, :k]]
    return voxels, coors, num
