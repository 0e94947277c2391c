"""pf_scatter_static -- the allocation-free BEV scatter (include/pillarfront.h) -- against pf_scatter, both through the
C ABI.  The entry point has no counterpart in the reference, so the
existing entry point is the yardstick: the canvas is copy work, compared BIT-EXACT like every other scatter test.

Collected last on purpose (file name).  Passes alone in a fresh process and after
the other tests: it only assumes a CUDA device and resets what it uses of the shared static cell map itself.
"""
import importlib

import numpy as np
import pytest

from tests.util import cfgmod

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

N_CH = 64


def _case(rng, nx, ny, batch, per_frame, dev):
    """Random pillars: unique cells per frame plus a few duplicated cells (highest row wins), frame id in column 3."""
    coors = []
    for b, v in enumerate(per_frame):
        cells = rng.choice(nx * ny, size=v, replace=False)
        if v > 20:
            cells[-10:] = cells[:10]                      # duplicates: the later row must win
        coors.append(np.stack([cells // ny, cells % ny, np.zeros(v, np.int64), np.full(v, b)], 1).astype(np.int32))
    coors = np.concatenate(coors)
    coors[::97, 0] = -1                                   # out-of-range rows are ignored by both entry points
    feats = rng.standard_normal((len(coors), N_CH)).astype(np.float32)
    return torch.from_numpy(feats).to(dev), torch.from_numpy(coors).to(dev)


def _both(lib_mod, handle, feats, coors, batch, nx, ny, out_static, out_handle):
    L = handle.lib
    stream = torch.cuda.current_stream().cuda_stream
    n = int(coors.shape[0])
    out_static.fill_(7.0)                                 # every cell must be rewritten
    out_handle.fill_(-7.0)
    lib_mod.check(L.pf_scatter_static(feats.data_ptr(), N_CH, 1, coors.data_ptr(), int(coors.shape[1]), n, batch, N_CH,
                                      nx, ny, out_static.data_ptr(), stream), "pf_scatter_static")
    lib_mod.check(L.pf_scatter(handle.ptr, feats.data_ptr(), N_CH, 1, coors.data_ptr(), int(coors.shape[1]), n, batch,
                               N_CH, out_handle.data_ptr(), stream), "pf_scatter")
    torch.cuda.synchronize()


def test_static_scatter_matches_pf_scatter_twice_in_a_row(pf):
    lib_mod = importlib.import_module(pf.__name__ + "._lib")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    cfg = cfgmod.builtin_config("20cm")
    geom = cfgmod.derive_grid(cfg)
    nx, ny = int(geom.grid_size[0]), int(geom.grid_size[1])
    batch = 2
    assert batch * nx * ny <= 1 << 22
    handle = lib_mod.Handle(geom, cfg["max_voxels"], cfg["max_num_points"], 4, batch, 4096, 0)
    rng = np.random.default_rng(20261105)
    a_static = torch.empty((batch, N_CH, nx, ny), device=dev)
    a_handle = torch.empty_like(a_static)
    first = _case(rng, nx, ny, batch, [9000, 4000], dev)
    second = _case(rng, nx, ny, batch, [300, 0], dev)     # far fewer pillars: nothing of call 1 may survive in the map
    _both(lib_mod, handle, *first, batch, nx, ny, a_static, a_handle)
    assert torch.equal(a_static.view(torch.int32), a_handle.view(torch.int32)), "call 1 differs from pf_scatter"
    assert int((a_static != 0).sum()) > 0
    free_before = torch.cuda.mem_get_info(dev)[0]
    _both(lib_mod, handle, *second, batch, nx, ny, a_static, a_handle)
    free_after = torch.cuda.mem_get_info(dev)[0]
    assert torch.equal(a_static.view(torch.int32), a_handle.view(torch.int32)), "call 2 differs from pf_scatter"
    assert free_after == free_before, "device memory was allocated by a call that must not allocate"
    occupied = (a_static[:, 0] != 0).sum().item()
    assert 0 < occupied <= 300, "stale cells of the first call leaked into the second"
    handle.close()

    # a different canvas shape right after (the shared array still holds the 20 cm layout), n_voxels == 0, and the limit
    cfg2 = cfgmod.builtin_config("nuscene")
    geom2 = cfgmod.derive_grid(cfg2)
    nx2, ny2 = int(geom2.grid_size[0]), int(geom2.grid_size[1])
    h2 = lib_mod.Handle(geom2, cfg2["max_voxels"], cfg2["max_num_points"], 4, 1, 4096, 0)
    b_static = torch.empty((1, N_CH, nx2, ny2), device=dev)
    b_handle = torch.empty_like(b_static)
    third = _case(rng, nx2, ny2, 1, [5000], dev)
    _both(lib_mod, h2, *third, 1, nx2, ny2, b_static, b_handle)
    assert torch.equal(b_static.view(torch.int32), b_handle.view(torch.int32)), "other shape differs from pf_scatter"
    empty = (third[0][:0], third[1][:0])
    _both(lib_mod, h2, *empty, 1, nx2, ny2, b_static, b_handle)
    assert torch.equal(b_static.view(torch.int32), b_handle.view(torch.int32)) and not b_static.any()
    h2.close()

    L = lib_mod.load()
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.pf_scatter_static(first[0].data_ptr(), N_CH, 1, first[1].data_ptr(), 4, 10, 7, N_CH, nx, ny,
                             a_static.data_ptr(), stream)                      # 7 x 800 x 800 > 2^22 cells
    assert rc == -2, "beyond the static maximum must be PF_ERR_CAPACITY"
    msg = L.pf_last_error_string().decode()
    assert "4194304" in msg and "pf_scatter_static" in msg
    rc = L.pf_scatter_static(first[0].data_ptr(), N_CH, 1, first[1].data_ptr(), 4, 10, 1, N_CH, 1 << 16, 1 << 16,
                             a_static.data_ptr(), stream)                      # nx * ny overflows int32
    assert rc == -2
    torch.cuda.synchronize()
