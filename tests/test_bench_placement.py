"""Multi-GPU host logic on CPU: bench.py's session→GPU placement and its world_size-2 gloo
reductions (no data-path collective exists to test: sessions share nothing)."""
import os
import socket
import subprocess
import sys
import textwrap

import pytest

import bench as sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_round_robin_placement():
    assert sharding.placement(8, 8) == [[i] for i in range(8)]
    assert sharding.placement(5, 2) == [[0, 2, 4], [1, 3]]
    assert sharding.sessions_of_rank(1, 4, 10) == [1, 5, 9]
    assert sorted(sum(sharding.placement(37, 8), [])) == list(range(37))
    with pytest.raises(ValueError):
        sharding.gpu_for_session(0, 0)
    with pytest.raises(ValueError):
        sharding.sessions_of_rank(2, 2, 4)


def test_aggregate_is_total_frames_over_slowest_rank():
    assert sharding.aggregate_fps([100, 100], [1.0, 2.0]) == 100.0
    with pytest.raises(ValueError):
        sharding.aggregate_fps([1], [1.0, 2.0])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_world_size_2_gloo_max_over_ranks(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {ROOT!r})
        import torch.distributed as dist
        import bench as sharding
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        dist.init_process_group("gloo", rank=rank, world_size=world)
        mine = sharding.sessions_of_rank(rank, world, 5)
        dist.barrier()
        t = sharding.reduce_over_ranks(dist, [1.0 + rank], "max")[0]          # rank 1 is the slow one
        assert sharding.reduce_over_ranks(dist, [1.0], "sum")[0] == world
        assert sharding.reduce_over_ranks(None, [3.5], "min") == [3.5]
        frames = [len(sharding.sessions_of_rank(r, world, 5)) * 10 for r in range(world)]
        fps = sharding.aggregate_fps(frames, [t] * world)
        dist.barrier()
        if rank == 0:
            print("RESULT", mine, t, fps)
        dist.destroy_process_group()
    """))
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=180) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    line = [l for l in outs[0][0].splitlines() if l.startswith("RESULT")][0]
    assert "[0, 2, 4] 2.0 25.0" in line
