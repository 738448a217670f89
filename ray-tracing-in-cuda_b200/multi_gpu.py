"""Partitioning of one image (or one animation) over the GPUs of a box — SURVEY.md §8(e).

The reference's only multi-GPU mechanism is "one process per GPU, one frame each" (gpu-version/blue.py:23-32,
blue2.py:23-32: eight `CUDA_VISIBLE_DEVICES=k ./build/parallel_compute -f frame.json &` then `wait`).  Here:
  * still images: the SAMPLES of every pixel are split over the ranks (Philox is keyed on the absolute sample
    index, so the sample set does not depend on the rank count), every rank accumulates float RGB sums for its
    slice, and ONE reduce(sum) to rank 0 over NCCL/NVLink (gloo in the CPU tests) finishes the image;
  * animations: frames are dealt round-robin (frame f -> rank f % world), no collective at all.
`torch.distributed` is plumbing only; the tensor handed to it is a zero-copy view of the library's device buffer.
"""


def spp_slice(spp, rank, world):
    """[begin, end) of the absolute sample indices rank `rank` renders; slices tile [0, spp) exactly."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    return rank * spp // world, (rank + 1) * spp // world


def accum_view(renderer, device_index):
    """torch view (no copy) of the renderer's W*H*3 float accumulation buffer, for the NCCL reduce."""
    import torch
    ptr, n = renderer.accum_device_ptr()

    class _CAI:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}
    return torch.as_tensor(_CAI(), device=f"cuda:{device_index}")


def reduce_image(accum, dst=0, group=None):
    """Sum of the per-rank accumulation buffers on rank `dst` (one collective per image).  In place."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def gather_ordered(accum, dst=0, group=None):
    """Bit-reproducible variant: rank `dst` receives every rank's partial buffer and adds them in rank order
    ((p0 + p1) + p2 ...), independent of the collective's internal reduction order.  Returns the sum on dst."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return accum
    rank = dist.get_rank(group)
    parts = [torch.empty_like(accum) for _ in range(world)] if rank == dst else None
    dist.gather(accum, parts, dst=dst, group=group)
    if rank != dst:
        return None
    out = parts[0].clone()
    for p in parts[1:]:
        out += p
    return out


def frames_for_rank(n_frames, rank, world):
    """blue.py:23-32: frame f renders on GPU f % world."""
    return [f for f in range(n_frames) if f % world == rank]
