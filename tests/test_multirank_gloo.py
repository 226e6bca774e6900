"""N > 1 host path on CPU: two ranks over `gloo`, landmark shards as bench.py hands them to the ranks, the per-rank partial
sums combined with all_reduce exactly where the GPU path uses ncclAllReduce (cost; gradient of the replicated camera
blocks).  The per-rank arithmetic is the oracle's -- this test is about the sharding and the exchange, not the kernels."""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from multiview_b200 import _abi, synth
import oracle.binding  # noqa: F401  (spawned workers do not go through conftest.py)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASE = dict(sensors="rig3", num_ref=12, num_landmarks=500, obs_per_track=5, outlier_frac=0.02)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
    pr = synth.make_problem(seed=8, **CASE)
    q = pr.shard(rank, world)
    b = _abi.Binding(lib, "oracle_", q)
    cost, r = b.evaluate(True)
    # Jacobian columns of the replicated blocks (beg pose 7 | end pose 7 | rig 7 of every CAM block): J'r summed per camera image
    res, J = b.evaluate_jacobian(0, 0, q.num_obs, mode=1)
    g = np.zeros((len(pr.cam_sensor), 21))
    np.add.at(g, q.cid, np.einsum("nrc,nr->nc", J[:, :, :21], res))
    b.close()
    buf = torch.tensor(np.concatenate([[cost, float(len(r)), float(q.num_landmarks)], g.ravel()]))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    lo, hi = q.truth["pid_range"]
    ranges = [None] * world
    dist.all_gather_object(ranges, (lo, hi))
    if rank == 0:
        np.save(out, buf.numpy())
        assert ranges[0][0] == 0 and ranges[-1][1] == pr.num_landmarks and all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
    dist.destroy_process_group()


def test_two_ranks_sum_to_the_unsharded_problem(tmp_path, oracle_lib):
    world = 2
    out = str(tmp_path / "reduced.npy")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    pr = synth.make_problem(seed=8, **CASE)
    b = _abi.Binding(oracle_lib, "oracle_", pr)
    cost, r = b.evaluate(True)
    res, J = b.evaluate_jacobian(0, 0, pr.num_obs, mode=1)
    g = np.zeros((len(pr.cam_sensor), 21))
    np.add.at(g, pr.cid, np.einsum("nrc,nr->nc", J[:, :, :21], res))
    b.close()
    assert abs(got[0] - cost) <= 1e-12 * cost
    assert got[1] == len(r) and got[2] == pr.num_landmarks
    assert np.abs(got[3:] - g.ravel()).max() <= 1e-10 * np.abs(g).max()
