"""Multi-GPU partitioning of the render path: sample-index ranges per rank and one
sum-reduce of the accumulated ImageBlock — the multi-process form of `result.put(block)`
(reference src/block.cpp:105-114, called under a mutex from src/main.cpp:116).

Camera samples are independent (PER_SAMPLE sampler streams), so the union of the ranks' ranges is
the same set of samples for any world size; scene and BVH are replicated per GPU."""
from __future__ import annotations


def sample_range(rank: int, world: int, spp: int, first: int = 0):
    """Contiguous, balanced split of sample indices [first, first + spp) over `world` ranks."""
    base, extra = divmod(spp, world)
    a = first + rank * base + min(rank, extra)
    return a, a + base + (1 if rank < extra else 0)


def reduce_block(block_tensor, dst: int = 0):
    """Sum the ranks' RGBW accumulators onto `dst` (NCCL over NVLink on GPUs, gloo in CPU tests)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(block_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return block_tensor
