#!/usr/bin/env python3
"""torchrun helper: how long does the library's gather + fold take on the root after its own chunks are rendered?  (N ranks)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import __graft_entry__ as ge, bench
rt = ge.load_package()
from rt_b200 import multi_gpu
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
stream = torch.cuda.Stream()
with torch.cuda.stream(stream):
    r = rt.Renderer(local, stream=stream.cuda_stream)
    multi_gpu.init_comm(r)
    upload, _ = bench.make_upload(rt, "rtiow_1080p_1024spp")
    upload(r)
    for _ in range(3):
        r.render_shard(seed=0x2023, spp_count=1024); r.gather_image(0)
    torch.cuda.synchronize(); dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    res = []
    for k in range(5):
        dist.barrier(); torch.cuda.synchronize()
        ev[0].record(stream); r.render_shard(seed=0x2023, spp_count=1024); ev[1].record(stream); r.gather_image(0); ev[2].record(stream)
        torch.cuda.synchronize()
        res.append((ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])))
    t = torch.tensor(res[1:], device=f"cuda:{local}").mean(0)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    if rank == 0:
        print("render ms per rank:", [round(float(x[0]), 2) for x in allt])
        print("gather+fold ms per rank (the root's includes waiting for the slowest sender):", [round(float(x[1]), 2) for x in allt])
r.close()
dist.destroy_process_group()
