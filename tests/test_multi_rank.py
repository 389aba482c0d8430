"""world_size-2 gloo checks (CPU) of the N > 1 plumbing bench.py uses: utterances shard without overlap, the
timed region is bracketed by a barrier, durations reduce with MAX over ranks, whole-job rate is weak-scaling, and
the reference arm does its work on rank 0 only."""
import json
import os
import socket
import subprocess
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from chatterbox_b200 import dispatch as D

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = D.shard(7, rank, world)
    D.barrier(world)
    ms = D.max_over_ranks([10.0 + 5.0 * rank, 3.0 - rank], "cpu", world)     # rank 1 is the slow one on [0]
    D.barrier(world)
    q.put((rank, mine, ms))
    dist.destroy_process_group()


def test_two_ranks_shard_and_reduce():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, s0, m0), (r1, s1, m1) = got
    assert sorted(s0 + s1) == list(range(7)) and not set(s0) & set(s1)       # a partition of the utterances
    assert s0 == [0, 2, 4, 6] and s1 == [1, 3, 5]
    assert m0 == m1 == [15.0, 3.0]                                           # MAX over ranks, on every rank
    assert D.whole_job_rate(256 * 5, world, 1000.0) == 2560.0               # weak scaling: units x world / slowest


def test_single_rank_is_a_no_op():
    assert D.shard(3, 0, 1) == [0, 1, 2]
    assert D.max_over_ranks([1.5], "cpu", 1) == [1.5]
    D.barrier(1)


def test_reference_arm_runs_on_rank_zero_only():
    """`bench.py --impl reference` under torchrun: ranks > 0 exit 0 without work or output."""
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
