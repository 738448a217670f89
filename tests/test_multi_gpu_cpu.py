"""World-size-2 and -3 tests of the multi-GPU host logic on CPU (gloo): the sample chunks of an image are dealt to the ranks
(rt_rank_chunks / rt_sample_chunk and their Python mirrors), each rank "renders" its chunks (the CPU oracle standing in for a
GPU), the chunk sums travel to rank 0 and are added in chunk order: BIT-IDENTICAL to the single-rank image, whatever the rank
count.  (On the GPU the transfers are NCCL send/receive inside the library, csrc/rt_multi.cu; the -m gpu tests cover those.)"""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import __graft_entry__ as ge

ge.load_package()     # spawned workers import this module without conftest.py
import orc  # noqa: E402
from rt_b200 import animation, capi, multi_gpu  # noqa: E402


def test_chunk_dealing_rules_match_the_library():
    lib = capi.load_library()
    a, b = C.c_int(), C.c_int()
    for spp in (1, 5, 7, 8, 10, 1024, 4096):
        bounds = []
        for k in range(8):
            lib.rt_sample_chunk(spp, k, C.byref(a), C.byref(b))
            assert (a.value, b.value) == multi_gpu.sample_chunk(spp, k)
            bounds.append((a.value, b.value))
        assert bounds[0][0] == 0 and bounds[-1][1] == spp and all(x[1] == y[0] for x, y in zip(bounds, bounds[1:]))
        for world in (1, 2, 3, 4, 5, 8):
            sl = [multi_gpu.spp_slice(spp, r, world) for r in range(world)]
            assert sl[0][0] == 0 and sl[-1][1] == spp and all(x[1] == y[0] for x, y in zip(sl, sl[1:]))
            for r in range(world):
                lib.rt_rank_chunks(world, r, C.byref(a), C.byref(b))
                assert (a.value, b.value) == multi_gpu.rank_chunks(world, r)
    with pytest.raises(ValueError):
        multi_gpu.rank_chunks(2, 2)
    assert multi_gpu.frames_for_rank(20, 3, 8) == animation.frames_for_rank(20, 3, 8) == [3, 11, 19]
    assert sorted(sum((multi_gpu.frames_for_rank(180, r, 8) for r in range(8)), [])) == list(range(180))


W, H, SPP = 48, 27, 21


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = orc.load_scene(os.path.join(orc.GOLDEN, "sample_scene.json"), width=W, height=H, spp=SPP)
    k0, k1 = multi_gpu.rank_chunks(world, rank)
    mine = {}
    for k in range(k0, k1):                       # one buffer per chunk, each a sequential sum over the chunk's samples
        b, e = multi_gpu.sample_chunk(SPP, k)
        if e > b:
            mine[k] = orc.render_range(sc, 11, b, e - b, SPP, threads=1)[0]
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0)
    if rank == 0:
        chunks = {}
        for part in gathered:
            chunks.update(part)
        np.save(os.path.join(out_dir, f"image_{world}.npy"), multi_gpu.fold_chunks(chunks))
    # the whole slice of a rank in one oracle call folds the same chunks in the same order
    s0, s1 = multi_gpu.spp_slice(SPP, rank, world)
    if s1 > s0:
        whole = orc.render_range(sc, 11, s0, s1 - s0, SPP, threads=1)[0]
        assert np.array_equal(whole, multi_gpu.fold_chunks(mine))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_fold_to_the_single_rank_image_bit_for_bit(tmp_path, world):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sc = orc.load_scene(os.path.join(orc.GOLDEN, "sample_scene.json"), width=W, height=H, spp=SPP)
    full, _ = orc.render(sc, seed=11, threads=1)
    assert np.array_equal(np.load(tmp_path / f"image_{world}.npy"), full)
