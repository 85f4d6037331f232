"""Placement of independent desktop sessions on the GPUs of one box.

The reference's only multi-GPU notion is device selection: `--gpu_id`
(src/selkies/legacy/webrtc.py:459-461) → `cuda-device-id` on the stage and the
matching NVENC factory (gstwebrtc_app.py:264-272, :289-298).  Sessions share
nothing, so there is no collective: a rank owns its sessions outright, and the
only cross-rank traffic in bench.py is the barrier and the max of the timings.
"""
from __future__ import annotations


def gpu_for_session(session_index: int, n_gpus: int) -> int:
    """Round-robin: session i → GPU i mod N."""
    if n_gpus < 1:
        raise ValueError("n_gpus must be >= 1")
    if session_index < 0:
        raise ValueError("session_index must be >= 0")
    return session_index % n_gpus


def sessions_of_rank(rank: int, world_size: int, total_sessions: int) -> list[int]:
    """Global session ids owned by `rank` (one rank per GPU)."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    return [i for i in range(total_sessions) if gpu_for_session(i, world_size) == rank]


def placement(total_sessions: int, n_gpus: int) -> list[list[int]]:
    return [sessions_of_rank(r, n_gpus, total_sessions) for r in range(n_gpus)]


def aggregate_fps(frames_per_rank: list[int], seconds_per_rank: list[float]) -> float:
    """Whole-job throughput: all frames over the slowest rank's time (weak scaling)."""
    if not frames_per_rank or len(frames_per_rank) != len(seconds_per_rank):
        raise ValueError("one frame count and one time per rank")
    return sum(frames_per_rank) / max(seconds_per_rank)


def reduce_max_time(dist, seconds: float) -> float:
    """max over ranks of a host float through torch.distributed (gloo or nccl-free)."""
    import torch
    t = torch.tensor([seconds], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])
