"""Sequence-level sharding across the GPUs of one box (reference: tools/eval.py:137-143 spawns one
process per GPU; networks/managers/evaluator.py:276-295 hands out sequence indices through a queue).

Engines share nothing but read-only weights, so the hot path needs no collective: sequence i runs on
rank i mod world.  The only exchange is one final gather of the per-rank results (frames, seconds,
mask checksums) with ``torch.distributed.all_gather_object`` -- NCCL on the GPU box, gloo in the CPU
tests."""
import torch.distributed as dist


def assign_sequences(n_sequences, world_size, rank):
    """Static round-robin partition (equal-length synthetic clips; same rule as i % gpu_num)."""
    if not 0 <= rank < world_size:
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    return list(range(rank, n_sequences, world_size))


def gather_results(local_results):
    """local_results: {sequence_id: result}.  Returns the merged dict on every rank and checks that no
    sequence was processed twice."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return dict(local_results)
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, dict(local_results))
    merged = {}
    for part in parts:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError("sequence %r was processed by two ranks" % (k,))
            merged[k] = v
    return merged


def max_over_ranks(value, device=None):
    """Max of a scalar over ranks (device timings are reported as the max over ranks)."""
    import torch
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])
