"""Multi-GPU form of the path: frames (or stereo pairs) are independent units,
frame i belongs to rank i % world; there is no data-path collective.  The only communication is
the gather of the fixed-capacity result slabs {header, keypoints, descriptors} (NCCL over NVLink on
the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def frame_owner(i: int, world: int) -> int:
    return i % world


def local_frames(n_frames: int, world: int, rank: int):
    """Global frame indices processed by `rank`, in local order."""
    return list(range(rank, n_frames, world))


def gather_slabs(slabs, world: int, group=None):
    """all_gather of equally-sized per-rank tensors; returns tensors of shape [world, *local]."""
    out = []
    for t in slabs:
        g = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        if world == 1:
            g[0].copy_(t)
        else:
            dist.all_gather_into_tensor(g.view(-1), t.contiguous().view(-1), group=group)
        out.append(g)
    return out


def to_global_order(g: torch.Tensor, n_frames: int):
    """g: [world, n_local, ...] gathered slabs -> [n_frames, ...] in global frame order
    (frame i = rank i % world, local index i // world)."""
    world, n_local = g.shape[0], g.shape[1]
    inter = g.transpose(0, 1).reshape((world * n_local,) + tuple(g.shape[2:]))
    return inter[:n_frames]
