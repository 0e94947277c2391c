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


_CONCAT_KEYS = frozenset({"voxels", "num_points_per_voxel", "match_indices_num"})   # joined along axis 0
_INT32_KEYS = frozenset({"coordinates", "num_points_per_voxel", "voxel_num"})


def _stack(elems):
    if any(_is_tensor(e) for e in elems):
        import torch
        dev = next(e.device for e in elems if _is_tensor(e))
        return torch.stack([e if _is_tensor(e) else torch.as_tensor(e, device=dev) for e in elems], dim=0)
    return np.stack(elems, axis=0)


def merge_second_batch(batch_list, _unused=False):
    """Collate per-frame example dicts (framework/utils.py:23-48): see the module docstring for the layout."""
    columns = defaultdict(list)
    for example in batch_list:
        for key, value in example.items():
            columns[key].append(value)
    pad_frame = len(batch_list) > 1            # the reference pads the frame id only for a real batch
    merged = {}
    for key, elems in columns.items():
        if key == "coordinates":
            merged[key] = _cat([_pad_frame_id(c, i) for i, c in enumerate(elems)] if pad_frame else elems)
        elif key in _CONCAT_KEYS:
            merged[key] = _cat(elems)
        else:
            merged[key] = _stack(elems)
    return merged


def example_convert_to_torch(example, dtype=None, device=None):
    """Move the tensors of an example to ``device`` (framework/utils.py:7-20): ``voxels`` in ``dtype`` (float32),
    coordinates / counts as int32, ``anchors_mask`` as bool, everything else untouched.  Tensors already there with
    the right dtype are returned as they are (no copy)."""
    import torch
    device = torch.device("cuda:0") if device is None else torch.device(device)
    casts = {"voxels": torch.float32 if dtype is None else dtype, "anchors_mask": torch.bool}
    casts.update({k: torch.int32 for k in _INT32_KEYS})
    out, pending = {}, False
    for k, v in example.items():
        if k not in casts:
            out[k] = v
        elif device.type == "cuda" and isinstance(v, np.ndarray) and v.flags.c_contiguous and v.flags.writeable:
            # numpy -> CUDA: all copies are issued first and waited for ONCE (torch.as_tensor waits after each).  The arrays
            # VoxelGenerator.generate returns are views of pinned blocks, so these are plain DMA transfers; the single wait
            # below keeps the reference's contract that the source may be dropped as soon as this function returns.
            t = torch.from_numpy(v).to(device, non_blocking=True)
            out[k] = t if t.dtype == casts[k] else t.to(casts[k])
            pending = True
        else:
            out[k] = torch.as_tensor(v, dtype=casts[k], device=device)
    if pending:
        torch.cuda.current_stream(device).synchronize()
    return out
