"""Partitioning of one image (or one animation) over the GPUs of a box — SURVEY.md §8(e).

The reference's only multi-GPU mechanism is "one process per GPU, one frame each" (gpu-version/blue.py:23-32,
blue2.py:23-32: eight `CUDA_VISIBLE_DEVICES=k ./build/parallel_compute -f frame.json &` then `wait`).  Here:
  * still images: the samples of every pixel fall into RT_SAMPLE_CHUNKS = 8 chunks (rt.h: rt_sample_chunk; Philox is keyed on the
    absolute sample index, so the sample set does not depend on the rank count); rank r renders the whole chunks
    rt_rank_chunks(world, r) into per-chunk float buffers, and the LIBRARY gathers them on the root over NCCL/NVLink and adds them
    in chunk order (csrc/rt_multi.cu) — bit-identical images for 1, 2, 4, 8 GPUs;
  * animations: frames are dealt round-robin (frame f -> rank f % world), no collective at all.
This module mirrors the dealing rules in Python (tests, bench) and hands the NCCL unique id from rank 0 to the other ranks
through torch.distributed — plumbing only: the communicator, the transfers and the sum are the library's.
"""
RT_SAMPLE_CHUNKS = 8


def sample_chunk(total_spp, k):
    """[begin, end) of chunk k of an image of total_spp samples (rt_sample_chunk)."""
    return k * total_spp // RT_SAMPLE_CHUNKS, (k + 1) * total_spp // RT_SAMPLE_CHUNKS


def rank_chunks(world, rank):
    """chunks [k0, k1) of rank `rank` (rt_rank_chunks): contiguous, as even as eight chunks allow."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    return rank * RT_SAMPLE_CHUNKS // world, (rank + 1) * RT_SAMPLE_CHUNKS // world


def spp_slice(spp, rank, world):
    """[begin, end) of the absolute sample indices rank `rank` renders: its whole chunks; slices tile [0, spp) exactly."""
    k0, k1 = rank_chunks(world, rank)
    return sample_chunk(spp, k0)[0], (sample_chunk(spp, k1 - 1)[1] if k1 > k0 else sample_chunk(spp, k0)[0])


def fold_chunks(chunk_sums):
    """(((c0 + c1) + c2) + ...) over the non-empty chunks in ascending chunk order — what k_fold_chunks does on the root.
    chunk_sums: {chunk index: float32 array}."""
    out = None
    for k in sorted(chunk_sums):
        out = chunk_sums[k].copy() if out is None else out + chunk_sums[k]
    return out


def init_comm(renderer, group=None):
    """Gives `renderer` (one per rank) the library's NCCL communicator: rank 0 draws the unique id, torch.distributed carries
    its 128 bytes to the other ranks (any backend), every rank calls rt_comm_init."""
    import torch.distributed as dist
    world, rank = (dist.get_world_size(group), dist.get_rank(group)) if dist.is_initialized() else (1, 0)
    if world == 1:
        return
    box = [type(renderer).comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    renderer.comm_init(box[0], world, rank)


def frames_for_rank(n_frames, rank, world):
    """blue.py:23-32: frame f renders on GPU f % world."""
    return [f for f in range(n_frames) if f % world == rank]
