"""CPU, world_size 2 over gloo: env-wise sharding and the one collective of the path (verify-statistics reduction).

The per-rank arithmetic here is the host build of the tick logic (tests/host_emul); what is under test is the host-side
multi-process plumbing of human_robot_gym_b200/sharded.py: partition, independence of the shards, SUM/MIN reduction."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from human_robot_gym_b200 import sharded
from tests import parity_util as pu


def test_shard_range_partitions_every_env_once():
    for n in (1, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 8):
            covered = []
            for r in range(world):
                lo, hi = sharded.shard_range(n, r, world)
                assert 0 <= lo <= hi <= n
                covered += list(range(lo, hi)) if n < 100 else [(lo, hi)]
            if n < 100:
                assert covered == list(range(n))
            else:
                assert covered[0][0] == 0 and covered[-1][1] == n
                assert all(a[1] == b[0] for a, b in zip(covered, covered[1:]))
                sizes = [b - a for a, b in covered]
                assert max(sizes) - min(sizes) <= 1
    assert sharded.shard_range(65536, 3, 8) == (24576, 32768)
    with pytest.raises(ValueError):
        sharded.shard_range(8, 8, 8)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run_shard(rank, world, port, n_total, ticks, seed, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from human_robot_gym_b200.cabi import make_cfg
        from human_robot_gym_b200.config import ShieldConfig
        from human_robot_gym_b200.synthetic import SyntheticScenario
        from tests.host_emul import emul
        files = pu.default_files("schunk")
        sc = ShieldConfig.from_files(0.004, *files)
        # every rank generates the GLOBAL synthetic inputs from the same seed and takes its rows
        scen = SyntheticScenario(n_total, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, seed)
        lo, hi = sharded.shard_range(n_total, rank, world)
        n = hi - lo
        em = emul.EmulBatch(make_cfg(sc, 1024), n, np.zeros((n, 6)), scen.init_q[lo:hi], 1)
        phase = np.arange(n_total) % 25
        counts = np.zeros(10, np.int64)
        min_d = np.inf
        last = None
        for k in range(ticks):
            t = k * 0.004
            em.human_measurement(scen.keypoints(t)[lo:hi], t)
            g = phase == (k % 25)
            if g.any():
                goal = scen.next_goal(g)
                em.new_goal(goal[lo:hi], mask=g[lo:hi])
            last = em.step(t)
            counts[0] += n
            counts[1] += int((last["is_safe"] == 0).sum())
            counts[2:5] += np.bincount(last["branch"], minlength=3)[:3]
            counts[5] += int((last["status"] & 1).astype(bool).sum())
            md = em.capsules()[2]
            md = md[np.isfinite(md).all(axis=1)]
            if md.size:
                min_d = min(min_d, float(md.min()))
        st = torch.from_numpy(counts.copy())
        st[9:10] = torch.tensor([min_d], dtype=torch.float64).view(torch.int64)
        red = sharded.reduce_stats(st)
        ms = sharded.max_over_ranks(10.0 + rank, "cpu")
        # gather the last setpoints of every shard on every rank (test only; the product never gathers)
        parts = [None] * world
        dist.all_gather_object(parts, (lo, hi, last["q"], last["is_safe"], last["branch"]))
        if rank == 0:
            np.savez(os.path.join(out_dir, "sharded.npz"), q=np.concatenate([p[2] for p in parts]),
                     is_safe=np.concatenate([p[3] for p in parts]), branch=np.concatenate([p[4] for p in parts]),
                     n_steps=red["n_steps"], n_unsafe=red["n_unsafe"], n_branch=np.array(red["n_branch"]),
                     n_replan=red["n_replan"], min_dist=red["min_dist"], ms=ms,
                     local=np.concatenate([counts[:6], [min_d]]))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_ranks_equal_one_rank(tmp_path):
    n_total, ticks, seed = 26, 60, 1003
    port = _free_port()
    mp.spawn(_run_shard, args=(2, port, n_total, ticks, seed, str(tmp_path)), nprocs=2, join=True)
    two = np.load(tmp_path / "sharded.npz")
    # the same job in one process, no process group (reduce_stats then returns the local values)
    os.makedirs(tmp_path / "one", exist_ok=True)
    port = _free_port()
    mp.spawn(_run_shard, args=(1, port, n_total, ticks, seed, str(tmp_path / "one")), nprocs=1, join=True)
    one = np.load(tmp_path / "one" / "sharded.npz")
    assert np.array_equal(one["q"], two["q"])                   # shards are independent: bit-identical setpoints
    assert np.array_equal(one["is_safe"], two["is_safe"]) and np.array_equal(one["branch"], two["branch"])
    for key in ("n_steps", "n_unsafe", "n_replan"):
        assert int(one[key]) == int(two[key])
    assert np.array_equal(one["n_branch"], two["n_branch"])
    assert float(one["min_dist"]) == float(two["min_dist"])     # MIN over ranks
    assert int(two["n_steps"]) == n_total * ticks
    assert int(two["local"][0]) == 13 * ticks                   # rank 0 alone only saw its half
    assert float(two["ms"]) == 11.0                             # max over ranks of the per-rank time
