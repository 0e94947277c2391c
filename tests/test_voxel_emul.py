"""The voxelisation kernel bodies (csrc/pf_voxel_kernels.cuh) run on host threads (tests/emul/) against the oracle.

The build container has no GPU; the GPU parity tests own the real thing.  What this holds is the kernel LOGIC: first
index per cell, the ticketed look-back scan, the cut-off index of the max_voxels `break`, in-tile ranks, the
(tile x voxel) count matrix with its partial rows and saturating bytes, the slot lists, and the self-cleaning grid
across calls (the emulated workspace starts as garbage on purpose)."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle
from tests.emul import build as emul_build
from tests.util import cfgmod, synth


@pytest.fixture(scope="module")
def emul():
    L = C.CDLL(str(emul_build.build_voxel()))
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
    L.emul_vox_create.restype = C.c_void_p
    L.emul_vox_create.argtypes = [fp, fp, ip, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    L.emul_vox_destroy.argtypes = [C.c_void_p]
    L.emul_voxelize.restype = C.c_int
    L.emul_voxelize.argtypes = [C.c_void_p, fp, ip, C.c_int, C.c_int, C.c_int, C.c_int, fp, ip, ip, ip]
    L.emul_consume_map.restype = C.c_int
    L.emul_consume_map.argtypes = [C.c_void_p]
    return L


class Emul:
    def __init__(self, L, geom, max_voxels, P, F, max_frames, cap):
        self.L, self.mv, self.P, self.F = L, max_voxels, P, F
        off = np.ascontiguousarray(geom.offset, np.float32)
        vs = np.ascontiguousarray(geom.voxel_size, np.float32)
        gs = np.ascontiguousarray(geom.grid_size, np.int32)
        fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
        self.h = L.emul_vox_create(off.ctypes.data_as(fp), vs.ctypes.data_as(fp), gs.ctypes.data_as(ip), max_voxels, P, F,
                                   max_frames, cap)

    def run(self, frames, max_frame_points=None, keep=False):
        fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
        B = len(frames)
        sizes = [len(f) for f in frames]
        pts = np.ascontiguousarray(np.concatenate(frames + [np.zeros((1, self.F), np.float32)]), np.float32)
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
        vox = np.full((B * self.mv, self.P, self.F), 7.0, np.float32)
        coors = np.full((B * self.mv, 3), -7, np.int32)
        num = np.full((B * self.mv,), -7, np.int32)
        vnum = np.full((B,), -7, np.int32)
        mfp = max(sizes + [1]) if max_frame_points is None else max_frame_points
        rc = self.L.emul_voxelize(self.h, pts.ctypes.data_as(fp), offs.ctypes.data_as(ip), B, mfp, self.F, 1 if keep else 0,
                                  vox.ctypes.data_as(fp), coors.ctypes.data_as(ip), num.ctypes.data_as(ip),
                                  vnum.ctypes.data_as(ip))
        assert rc == 0, f"workspace invariant broken after the call (code {rc})"
        out = []
        for f in range(B):
            sl = slice(f * self.mv, f * self.mv + int(vnum[f]))
            out.append((vox[sl], coors[sl], num[sl]))
        return out

    def close(self):
        self.L.emul_vox_destroy(self.h)


def _same(got, want, what):
    gv, gc, gn = got
    wv, wc, wn = want
    assert len(gn) == len(wn), f"{what}: {len(gn)} voxels, oracle {len(wn)}"
    assert np.array_equal(gn, wn), f"{what}: counts"
    assert np.array_equal(gc, wc), f"{what}: coors"
    assert np.array_equal(gv.view(np.uint32), np.ascontiguousarray(wv).view(np.uint32)), f"{what}: voxels"


@pytest.mark.parametrize("gkey,mv,P", [("20cm", 700, 15), ("nuscene", 300, 100), ("10cm", 5000, 3)])
def test_emulated_chain_matches_oracle(emul, gkey, mv, P):
    """Several tiles per frame, overflowing and non-overflowing frames, an empty and a one-point frame, two calls on the
    same workspace (tickets / status words / flags are per-call state; count matrix and grid clean themselves)."""
    cfg = cfgmod.builtin_config(gkey)
    geom = cfgmod.derive_grid(cfg)
    spec = [("S1", 9000, 500), ("S3", 5000, 0), ("S2", 6400, 0), ("S4", 4200, 300), ("S1", 0, 1), ("S1", 1, 1), ("S1", 7000, 12)]
    frames = []
    for i, (kind, n, cells) in enumerate(spec):
        f = synth.make_frame(kind, geom, max(n, 8), 600 + i, max(cells, 1))
        frames.append(np.ascontiguousarray(f[:n] if kind != "S4" else f, np.float32))
    em = Emul(emul, geom, mv, P, 4, len(frames), max(len(f) for f in frames))
    refs = [oracle.generate(geom, f, mv, P) for f in frames]
    assert any(len(r[2]) == mv for r in refs) and any(0 < len(r[2]) < mv for r in refs)
    for rep in range(2):
        got = em.run(frames if rep == 0 else frames[::-1])
        for f, (g, r) in enumerate(zip(got, refs if rep == 0 else refs[::-1])):
            _same(g, r, f"{gkey} rep {rep} frame {f}")
    em.close()


def test_emulated_chain_3d_grid_and_feature_counts(emul):
    rng = np.random.default_rng(77)
    for case in range(6):
        nx, ny, nz = int(rng.integers(2, 30)), int(rng.integers(2, 30)), int(rng.integers(1, 5))
        F = int(rng.integers(3, 7))
        vs = np.array([0.5, 0.25, 1.0])
        lo = np.array([-3.0, 1.0, -2.0])
        cfg = dict(detection_range=list(lo) + list(lo + vs * [nx, ny, nz]), voxel_size=list(vs),
                   max_voxels=int(rng.integers(1, 300)), max_num_points=int(rng.integers(1, 40)), num_point_features=F)
        geom = cfgmod.derive_grid(dict(cfg))
        n = int(rng.choice([1, 300, 2048, 2049, 6000]))
        span = vs * [nx, ny, nz]
        pts = (lo - 0.2 * span + rng.random((n, 3)) * 1.4 * span).astype(np.float32)
        pts = np.ascontiguousarray(np.concatenate([pts, rng.random((n, F - 3), dtype=np.float32)], axis=1))
        if n > 8:
            pts[rng.integers(0, n, n // 4)] = pts[rng.integers(0, n, n // 4)]          # duplicates
        em = Emul(emul, geom, cfg["max_voxels"], cfg["max_num_points"], F, 1, n)
        _same(em.run([pts])[0], oracle.generate(geom, pts, cfg["max_voxels"], cfg["max_num_points"]), f"case {case} {cfg}")
        em.close()


def test_emulated_saturating_counters_and_kept_map(emul):
    """One cell holding > 255 points per tile and P = 255 (byte counters saturate); the kept cell map is what the
    canvas fill consumes: one negative entry per voxel."""
    cfg = cfgmod.builtin_config("20cm")
    geom = cfgmod.derive_grid(cfg)
    f = synth.make_frame("S1", geom, 7000, 5, 9)          # 9 cells, ~780 points each
    for P in (255, 100, 1):
        em = Emul(emul, geom, 50, P, 4, 1, 7000)
        want = oracle.generate(geom, f, 50, P)
        got = em.run([f], keep=True)[0]
        _same(got, want, f"P={P}")
        assert emul.emul_consume_map(em.h) == len(want[2])
        _same(em.run([f])[0], want, f"P={P} second call")
        em.close()


def test_emulated_overlong_frame_is_truncated(emul):
    """a frame longer than the call's max_frame_points is processed up to that bound, deterministically."""
    cfg = cfgmod.builtin_config("20cm")
    geom = cfgmod.derive_grid(cfg)
    f = synth.make_frame("S3", geom, 6000, 8)
    em = Emul(emul, geom, 4000, 15, 4, 2, 6000)
    got = em.run([f, f[:3000]], max_frame_points=4096)
    _same(got[0], oracle.generate(geom, f[:4096], 4000, 15), "truncated")
    _same(got[1], oracle.generate(geom, f[:3000], 4000, 15), "short")
    em.close()


def test_emulated_many_tiles_per_frame(emul):
    """17 tiles per frame: a voxel whose points are spread over every tile (carry down the whole column of the count
    matrix), with saturating P = 255 and with an overflowing max_voxels."""
    cfg = cfgmod.builtin_config("20cm")
    geom = cfgmod.derive_grid(cfg)
    f = synth.make_frame("S1", geom, 34000, 3, 40)           # 40 cells, ~800 points each, in every tile
    em = Emul(emul, geom, 64, 255, 4, 1, 34000)
    _same(em.run([f])[0], oracle.generate(geom, f, 64, 255), "17 tiles, P = 255")
    em.close()
    em = Emul(emul, geom, 30, 9, 4, 1, 34000)                # overflows max_voxels as well
    _same(em.run([f])[0], oracle.generate(geom, f, 30, 9), "17 tiles, overflow")
    em.close()
