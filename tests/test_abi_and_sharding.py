"""CPU-side checks: the C-ABI library loads and exports every symbol include/*.h declares (no
compute calls without a GPU), and the multi-GPU host logic (frame sharding + result gather
ordering) over a world_size-2 gloo group."""
import ctypes as C
import glob
import os
import re
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for hdr in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = re.sub(r"/\*.*?\*/", "", open(hdr).read(), flags=re.S)
        names |= set(re.findall(r"\b(orb[xm]_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    from orb_slam3_mac_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 19
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.orbx_abi_version() == 1
    for n in names:  # every declared symbol also has a ctypes prototype in the Python mirror
        if n.startswith("orbx_"):
            assert n in _lib.ORBX_SYMBOLS, n


def test_create_fails_loudly_without_gpu():
    """No CPU fallback: on a box without a CUDA device orbx_create must return ORBX_E_CUDA."""
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    from orb_slam3_mac_b200 import _lib
    from orb_slam3_mac_b200.extractor import ORBextractor
    with pytest.raises(_lib.OrbxError) as e:
        ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == _lib.ORBX_E_CUDA
    assert "no CPU fallback" in str(e.value)


def test_bad_config_is_rejected_before_touching_cuda():
    from orb_slam3_mac_b200 import _lib
    lib = _lib.load()
    h = C.c_void_p()
    for bad in (dict(nlevels=0), dict(nlevels=17), dict(nfeatures=0), dict(scale_factor=1.0), dict(max_batch=0)):
        kw = dict(nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th_fast=20, min_th_fast=7, max_width=752,
                  max_height=480, max_batch=1, n_slots=1, device=0, flags=0)
        kw.update(bad)
        cfg = _lib.OrbxConfig(**kw)
        assert lib.orbx_create(C.byref(cfg), C.byref(h)) == _lib.ORBX_E_BAD_ARG, bad
    assert lib.orbx_create(None, C.byref(h)) == _lib.ORBX_E_BAD_ARG


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from orb_slam3_mac_b200 import sharding
    mine = sharding.local_frames(n_frames, world, rank)
    n_local = (n_frames + world - 1) // world
    cap = 5
    # fake result slabs: header {n, mono, status, 0}, keypoint rows tagged with the global frame index
    hdr = torch.zeros((n_local, 4), dtype=torch.int32)
    kps = torch.zeros((n_local, cap, 7), dtype=torch.float32)
    for j, f in enumerate(mine):
        hdr[j, 0] = f % cap
        hdr[j, 1] = f
        kps[j, :, 0] = float(f)
    g_hdr, g_kps = sharding.gather_slabs([hdr, kps], world)
    gh = sharding.to_global_order(g_hdr, n_frames)
    gk = sharding.to_global_order(g_kps, n_frames)
    ok = bool((gh[:, 1] == torch.arange(n_frames, dtype=torch.int32)).all()) and \
        bool((gk[:, 0, 0] == torch.arange(n_frames, dtype=torch.float32)).all())
    ret[rank] = ok
    dist.destroy_process_group()


@pytest.mark.parametrize("n_frames", [8, 7])
def test_frame_sharding_and_gather_order_gloo_world2(n_frames):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    ctx = mp.get_context("spawn")
    ps = [ctx.Process(target=_worker, args=(r, world, port, n_frames, ret)) for r in range(world)]
    [p.start() for p in ps]
    [p.join(120) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    assert ret.get(0) and ret.get(1)


def test_sharding_partition_is_exact():
    from orb_slam3_mac_b200 import sharding
    for world in (1, 2, 4, 8):
        for n in (1, 7, 8, 100):
            seen = sorted(i for r in range(world) for i in sharding.local_frames(n, world, r))
            assert seen == list(range(n))
            for r in range(world):
                assert all(sharding.frame_owner(i, world) == r for i in sharding.local_frames(n, world, r))


def _sink_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from orb_slam3_mac_b200 import sharding
    # the gathering rank exports a (here: fake) 64-byte IPC handle, every rank must end up with it
    h = sharding.exchange_sink_handle(bytes(range(64)) if rank == 0 else None, src=0)
    B, slots = 6, 2
    # what the fused gather does on the GPU box, replayed with a gloo gather: rank r "writes" its batch of slot sl
    # to sink frames [first, first + B); the slab of a slot then reads like an all_gather output
    slab = torch.full((sharding.sink_frames(slots, world, B),), -1, dtype=torch.int64)
    mine = torch.tensor(sharding.local_frames(world * B, world, rank), dtype=torch.int64)
    parts = [torch.zeros(B, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(parts, mine)
    for sl in range(slots):
        for r in range(world):
            f0 = sharding.sink_first_frame(sl, r, world, B)
            slab[f0:f0 + B] = parts[r] + 1000 * sl
    ok = h == bytes(range(64))
    for sl in range(slots):
        region = slab[sl * world * B:(sl + 1) * world * B].view(world, B)
        ok = ok and bool((sharding.to_global_order(region, world * B) == torch.arange(world * B) + 1000 * sl).all())
    ret[rank] = ok
    dist.destroy_process_group()


def test_fused_gather_layout_and_handle_exchange_gloo_world2():
    """Host logic of the fused result gather (orbx_sink_*): the IPC handle reaches every rank and the sink layout
    [slot][rank][local frame] maps back to global frame order."""
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    ctx = mp.get_context("spawn")
    ps = [ctx.Process(target=_sink_worker, args=(r, world, port, ret)) for r in range(world)]
    [p.start() for p in ps]
    [p.join(120) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    assert ret.get(0) and ret.get(1)
