"""N > 1 host logic on CPU: two gloo ranks shard a short stream frame-parallel (frame f -> rank
f mod G), stitch their frames with the CPU oracle standing in for the GPU, and rank 0 gets every
panorama digest back in stream order.  No collective touches pixel data."""
import hashlib
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, out_path):
    sys.path.insert(0, ROOT)
    import oracle
    sharding = importlib.import_module("image-stitching_b200.sharding")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        H = [np.array([[1.02, 0.01, 3.0], [-0.01, 1.03, 2.0], [1e-5, 0, 1.0]]),
             np.array([[0.98, -0.02, 70.0], [0.015, 1.01, 1.0], [0, 1e-5, 1.0]])]
        mine = {}
        for f in sharding.frames_for_rank(n_frames, rank, world):
            srcs = [oracle.gen_frame(96, 64, i, f, oracle.NOISE) for i in range(2)]
            pano = oracle.stitch(srcs, H, (160, 72), nthreads=1)
            mine[f] = hashlib.sha256(pano.tobytes()).hexdigest()
        slowest = sharding.max_over_ranks(10.0 + rank, dist)
        ordered = sharding.resequence(mine, n_frames, dist)
        if rank == 0:
            want = []
            for f in range(n_frames):
                srcs = [oracle.gen_frame(96, 64, i, f, oracle.NOISE) for i in range(2)]
                want.append(hashlib.sha256(oracle.stitch(srcs, H, (160, 72), nthreads=1).tobytes()).hexdigest())
            ok = ordered == want and slowest == 10.0 + world - 1
            with open(out_path, "w") as fh:
                fh.write("ok" if ok else f"bad: {ordered} vs {want}, slowest={slowest}")
        else:
            assert ordered is None
    finally:
        dist.destroy_process_group()


def test_partition_is_exact_and_disjoint():
    sharding = importlib.import_module("image-stitching_b200.sharding")
    for world in (1, 2, 4, 8):
        for n in (0, 1, 7, 64):
            parts = [sharding.frames_for_rank(n, r, world) for r in range(world)]
            flat = sorted(f for p in parts for f in p)
            assert flat == list(range(n))
            assert all(sharding.owner_of_frame(f, world) == r for r, p in enumerate(parts) for f in p)
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_two_gloo_ranks_shard_and_resequence(tmp_path):
    out = tmp_path / "result.txt"
    mp.spawn(_worker, args=(2, _free_port(), 7, str(out)), nprocs=2, join=True)
    assert out.read_text() == "ok"


def test_resequence_rejects_wrong_owner():
    sharding = importlib.import_module("image-stitching_b200.sharding")
    assert sharding.resequence({0: "a", 1: "b"}, 2) == ["a", "b"]
    assert sharding.max_over_ranks(3.5) == 3.5
