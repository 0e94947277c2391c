"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Golden vector for the collate row (N3): the REFERENCE's own
merge_second_batch (framework/utils.py:23-48) on three seeded examples (one of them empty).

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_collate        (build container only: needs /root/reference)
"""
import numpy as np

from oracle.make_golden import GOLD, ref_modules


def examples():
    rng = np.random.default_rng(3)
    ex = []
    for v in (5, 0, 7):
        ex.append({"voxels": rng.random((v, 4, 4), dtype=np.float32),
                   "num_points_per_voxel": rng.integers(1, 5, v).astype(np.int32),
                   "coordinates": rng.integers(0, 50, (v, 3)).astype(np.int32),
                   "anchors_mask": rng.random(11) > 0.5, "match_indices_num": np.array([v], np.int32)})
    return ex


def main():
    ref_modules()
    from framework.utils import merge_second_batch
    ex = examples()
    out = {f"in{i}_{k}": v for i, e in enumerate(ex) for k, v in e.items()}
    for k, v in merge_second_batch(ex).items():
        out[f"merged_{k}"] = v
    for k, v in merge_second_batch(ex[:1]).items():
        out[f"single_{k}"] = v
    np.savez_compressed(GOLD / "collate.npz", **out)
    print({k: v.shape for k, v in out.items() if k.startswith(("merged", "single"))})


if __name__ == "__main__":
    main()
