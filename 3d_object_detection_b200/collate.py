"""Batch collation ("next" row N3): drop-ins for ``framework/utils.py:7-48`` that also accept the CUDA tensors
``VoxelGenerator.generate(..., device_outputs=True)`` returns, so a DataLoader ``collate_fn`` keeps working when the
front end's outputs stay on the device.

* ``merge_second_batch(batch_list)`` -- same keys, same layouts: ``voxels`` / ``num_points_per_voxel`` /
  ``match_indices_num`` concatenated along axis 0, ``coordinates`` concatenated with the frame id padded as the LAST
  column when the batch has more than one example (framework/utils.py:34-46), everything else stacked.  numpy
  examples give numpy (bit-for-bit the reference's result); CUDA examples are concatenated on their device.
* ``example_convert_to_torch(example, dtype, device)`` -- framework/utils.py:7-20; tensors already on ``device`` pass
  through without a copy.

For frames that are still raw point clouds, ``VoxelGenerator.generate_batch`` produces this collated layout
directly inside the voxelise kernels (compact rows, frame id column) with no concatenation at all.
"""
from __future__ import annotations

from collections import defaultdict

import numpy as np


def _is_tensor(x):
    return hasattr(x, "is_cuda") and hasattr(x, "dim")


def _cat(elems):
    if any(_is_tensor(e) for e in elems):
        import torch
        dev = next(e.device for e in elems if _is_tensor(e))
        return torch.cat([e if _is_tensor(e) else torch.as_tensor(e, device=dev) for e in elems], dim=0)
    return np.concatenate(elems, axis=0)


def _pad_frame_id(coor, i):
    if _is_tensor(coor):
        import torch
        return torch.nn.functional.pad(coor, (0, 1), mode="constant", value=i)
    return np.pad(coor, ((0, 0), (0, 1)), mode="constant", constant_values=i)


def merge_second_batch(batch_list, _unused=False):
    example_merged = defaultdict(list)
    for example in batch_list:
        for k, v in example.items():
            example_merged[k].append(v)
    ret = {}
    for key, elems in example_merged.items():
        if key in ("voxels", "num_points_per_voxel", "match_indices_num"):
            ret[key] = _cat(elems)
        elif key == "coordinates":
            if len(batch_list) > 1:
                ret[key] = _cat([_pad_frame_id(c, i) for i, c in enumerate(elems)])
            else:
                ret[key] = _cat(elems)
        elif any(_is_tensor(e) for e in elems):
            import torch
            dev = next(e.device for e in elems if _is_tensor(e))
            ret[key] = torch.stack([e if _is_tensor(e) else torch.as_tensor(e, device=dev) for e in elems], dim=0)
        else:
            ret[key] = np.stack(elems, axis=0)
    return ret


def example_convert_to_torch(example, dtype=None, device=None):
    import torch
    dtype = torch.float32 if dtype is None else dtype
    device = torch.device("cuda:0") if device is None else torch.device(device)
    example_torch = {}
    for k, v in example.items():
        if k in ("voxels",):
            example_torch[k] = torch.as_tensor(v, dtype=dtype, device=device)
        elif k in ("coordinates", "num_points_per_voxel", "voxel_num"):
            example_torch[k] = torch.as_tensor(v, dtype=torch.int32, device=device)
        elif k in ("anchors_mask",):
            example_torch[k] = torch.as_tensor(v, dtype=torch.bool, device=device)
        else:
            example_torch[k] = v
    return example_torch
