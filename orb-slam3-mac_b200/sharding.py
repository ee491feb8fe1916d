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


# ---- fused result gather (include/orbx.h, orbx_sink_*) ------------------------------------------------
# The product path for N > 1 does not call a collective at all: every rank's k_describe writes its result
# rows straight into ONE slab on the gathering rank (peer memory over NVLink).  The slab is laid out like
# the output of an all_gather -- [slot][rank][local frame] -- so to_global_order applies unchanged.
def sink_first_frame(slot: int, rank: int, world: int, batch: int) -> int:
    """First sink frame of `rank`'s batch in pipeline slot `slot`."""
    return (slot * world + rank) * batch


def sink_frames(n_slots: int, world: int, batch: int) -> int:
    return n_slots * world * batch


def exchange_sink_handle(handle: bytes | None, src: int = 0, group=None) -> bytes:
    """The gathering rank passes the 64-byte CUDA IPC handle of its sink (orbx_sink_export); every rank
    returns it.  Uses the process group the application already has (NCCL on the GPU box, gloo in tests)."""
    box = [handle if dist.get_rank(group) == src else None]
    dist.broadcast_object_list(box, src=src, group=group)
    h = box[0]
    if not isinstance(h, (bytes, bytearray)) or len(h) != 64:
        raise ValueError("sink handle must be 64 bytes")
    return bytes(h)


class ResultSink:
    """ctypes wrapper of orbx_sink: created on the gathering device, or opened from an exported handle."""

    def __init__(self, device: int, frames: int, cap: int, handle: bytes | None = None):
        import ctypes as C
        from . import _lib
        self._C, self._lib_mod = C, _lib
        self.lib = _lib.load()
        self.frames, self.cap, self.device = int(frames), int(cap), int(device)
        self.h = C.c_void_p()
        if handle is None:
            _lib.check(self.lib.orbx_sink_create(device, frames, cap, C.byref(self.h)))
        else:
            buf = (C.c_uint8 * 64).from_buffer_copy(handle)
            _lib.check(self.lib.orbx_sink_open(device, buf, frames, cap, C.byref(self.h)))

    def export(self) -> bytes:
        C = self._C
        buf = (C.c_uint8 * 64)()
        self._lib_mod.check(self.lib.orbx_sink_export(self.h, buf))
        return bytes(buf)

    def pointers(self):
        C = self._C
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._lib_mod.check(self.lib.orbx_sink_pointers(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def read_frame(self, frame: int):
        """-> (header[4] int32, keypoints[n], descriptors[n, 32]) of one gathered frame."""
        import numpy as np
        hdr = np.zeros(4, np.int32)
        self._lib_mod.check(self.lib.orbx_sink_read_frame(self.h, frame, hdr.ctypes.data, None, None, 0))
        n = int(min(max(hdr[0], 0), self.cap))
        kps = np.zeros(n, self._lib_mod.KP_DTYPE)
        desc = np.zeros((n, 32), np.uint8)
        if n:
            self._lib_mod.check(self.lib.orbx_sink_read_frame(self.h, frame, None, kps.ctypes.data, desc.ctypes.data, n))
        return hdr, kps, desc

    def close(self):
        if getattr(self, "h", None):
            self.lib.orbx_sink_destroy(self.h)
            self.h = None
