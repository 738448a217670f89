"""World-size-2 test of the multi-GPU host logic on CPU (gloo): sample slices tile the image's samples, each rank
"renders" its slice (with the CPU oracle standing in for a GPU), one reduce(sum) to rank 0 gives the same image as a
single-rank render; the ordered-gather variant is bit-identical to adding the slices in rank order."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import __graft_entry__ as ge

ge.load_package()     # spawned workers import this module without conftest.py
import orc  # noqa: E402
from rt_b200 import animation, multi_gpu  # noqa: E402


def test_spp_slices_tile_the_samples():
    for spp in (1, 7, 10, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            sl = [multi_gpu.spp_slice(spp, r, world) for r in range(world)]
            assert sl[0][0] == 0 and sl[-1][1] == spp
            assert all(a[1] == b[0] for a, b in zip(sl, sl[1:]))
    with pytest.raises(ValueError):
        multi_gpu.spp_slice(8, 2, 2)
    assert multi_gpu.frames_for_rank(20, 3, 8) == animation.frames_for_rank(20, 3, 8) == [3, 11, 19]
    assert sorted(sum((multi_gpu.frames_for_rank(180, r, 8) for r in range(8)), [])) == list(range(180))


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W, H, spp = 48, 27, 6
    sc = orc.load_scene(os.path.join(orc.GOLDEN, "sample_scene.json"), width=W, height=H, spp=spp)
    s0, s1 = multi_gpu.spp_slice(spp, rank, world)
    part, _ = orc.render(sc, seed=11, spp_begin=s0, spp_count=s1 - s0, threads=1)
    t = torch.from_numpy(part.copy()).reshape(-1)
    ordered = multi_gpu.gather_ordered(t.clone(), dst=0)
    multi_gpu.reduce_image(t, dst=0)
    if rank == 0:
        np.save(os.path.join(out_dir, "reduced.npy"), t.numpy().reshape(H, W, 3))
        np.save(os.path.join(out_dir, "ordered.npy"), ordered.numpy().reshape(H, W, 3))
    np.save(os.path.join(out_dir, f"part{rank}.npy"), part)
    dist.destroy_process_group()


def test_two_ranks_reduce_to_the_single_rank_image(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    W, H, spp = 48, 27, 6
    sc = orc.load_scene(os.path.join(orc.GOLDEN, "sample_scene.json"), width=W, height=H, spp=spp)
    full, _ = orc.render(sc, seed=11, threads=1)
    p0, p1 = np.load(tmp_path / "part0.npy"), np.load(tmp_path / "part1.npy")
    assert np.array_equal(np.load(tmp_path / "ordered.npy"), p0 + p1)             # fixed order: bit-exact
    assert np.array_equal(np.load(tmp_path / "reduced.npy"), p0 + p1)             # two ranks: one addition either way
    # same sample set as one rank; only the float summation order differs ((s0+s1+s2)+(s3+s4+s5) vs sequential)
    assert np.abs(np.load(tmp_path / "reduced.npy") - full).max() <= 1e-5 * spp
