"""The N > 1 path on CPU: two gloo ranks shard the requests, process their shard with the
oracle standing in for the device work, and reduce timings with MAX -- the same host logic
bench.py runs under torchrun with NCCL."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import kubectl_b200  # noqa: F401
from kubectl_b200 import dist_utils, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import py_oracle as po
    strings = synth.command_strings(total, seed=2)
    b, e = dist_utils.shard_range(total, rank, world)
    verdicts = [po.spec_verdict(s) for s in strings[b:e]]
    times = dist_utils.max_over_ranks([1.0 + rank, 5.0 - rank])
    counts = torch.tensor([e - b, sum(v == 0 for v in verdicts)], dtype=torch.int64)
    dist.all_reduce(counts)                      # test-only: the data path itself has no collective
    torch.save({"range": (b, e), "times": times, "counts": counts.tolist(), "verdicts": verdicts},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_two_ranks_cover_the_batch_and_reduce_with_max(tmp_path):
    total, world = 1001, 2
    mp.spawn(_worker, args=(world, _free_port(), total, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(tmp_path / f"r{i}.pt") for i in range(world)]
    assert r[0]["range"][0] == 0 and r[0]["range"][1] == r[1]["range"][0] and r[1]["range"][1] == total
    assert r[0]["times"] == r[1]["times"] == [2.0, 5.0]
    from oracle import py_oracle as po
    strings = synth.command_strings(total, seed=2)
    assert r[0]["verdicts"] + r[1]["verdicts"] == [po.spec_verdict(s) for s in strings]
    assert r[0]["counts"][0] == total


def test_shard_range_properties():
    for total in (0, 1, 7, 512, 513):
        for world in (1, 2, 3, 8):
            cuts = [dist_utils.shard_range(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in cuts]
            assert max(sizes) - min(sizes) <= 1
