"""Multi-GPU partitioning of the path: samples split over a replicated scene, ONE reduce of the film.

The reference has no distributed mode; its only multi-machine workflow renders independently and sums the EXR
accumulators offline with `master merge` (Options.cpp:1340-1409: out = a + b on all four channels, the fourth being
the sample count).  Here rank g of G renders the sample indices g, g + G, g + 2G, ... for every pixel — pixel-samples
are independent given the per-(pixel, sample) RNG streams — and the float64 (sum r, sum g, sum b, sum 1) films are
added with one all-reduce (NCCL over NVLink on the GPU box, gloo in the CPU tests).  No other communication.
"""
import torch
import torch.distributed as dist


def sample_split(rank, world, step, passes_per_step):
    """(first_sample, sample_stride) of `rank` for step `step`: global sample (step * passes + j) * world + rank."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return step * passes_per_step * world + rank, world


def samples_of(rank, world, step, passes_per_step):
    first, stride = sample_split(rank, world, step, passes_per_step)
    return [first + j * stride for j in range(passes_per_step)]


def reduce_film(accum, dst=None):
    """Sum the per-rank films in place: all ranks get the total (dst=None) or only `dst` does (merge_exr's a + b)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return accum
    if accum.dtype != torch.float64:
        raise TypeError("the film is a float64 (sum rgb, sum count) accumulator")
    if dst is None:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM)
    else:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum
