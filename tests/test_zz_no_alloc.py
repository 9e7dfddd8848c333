"""hrg_stage_reach_verify_noalloc (include/hrg_shield.h): the allocation-free form of the reach + verify stage, whose scratch
is a static __device__ array inside the library. No counterpart in the reference;
checked through the C ABI against the existing, allocating entry point on two calls in a row that use
different batch sizes and inputs, so stale contents of the shared array would show. Collected last on purpose (`-x`).
Does not need oracle/_ref: the existing entry point is the checker (it is tied to the reference by test_parity_gaps)."""
import ctypes as C

import numpy as np
import pytest

from tests import parity_util as pu

gpu = pytest.mark.gpu


def _inputs(sc, files, n, seed):
    from human_robot_gym_b200.synthetic import rest_pose
    J, K = sc.robot.nb_joints, len(pu.mocap_joint_names(files[2]))
    rng = np.random.default_rng(seed)
    pose = pu.random_base_pose(rng, n)
    q_lo, q_hi = np.asarray(sc.traj.q_min_allowed), np.asarray(sc.traj.q_max_allowed)
    q1 = rng.uniform(q_lo, q_hi, (n, J))
    q2 = np.clip(q1 + rng.uniform(-0.3, 0.3, (n, J)) * (rng.random((n, 1)) < 0.9), q_lo, q_hi)
    s_diff, t_reach = rng.uniform(0, 0.25, n), rng.uniform(0.004, 0.6, n)
    alpha = rng.uniform(0, 6, (n, J))
    kp = rest_pose(pu.mocap_joint_names(files[2]))[None] + pose[:, None, :3] + rng.uniform(-1.2, 1.2, (n, 1, 3)) * [1, 1, 0.3]
    kv = rng.uniform(-1.5, 1.5, (n, K, 3)) * (rng.random((n, 1, 1)) < 0.8)
    return pose, q1, q2, s_diff, t_reach, alpha, kp, kv


def _call(fn, name, cfg, n, ins, J, H):
    import torch

    from human_robot_gym_b200 import cabi
    outs = [torch.full((n, J, 7), np.nan, dtype=torch.float64, device="cuda:0"),
            torch.full((n, H, 7), np.nan, dtype=torch.float64, device="cuda:0"),
            torch.full((n, 3), np.nan, dtype=torch.float64, device="cuda:0"),
            torch.full((n,), 7, dtype=torch.uint8, device="cuda:0")]
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    cabi.check(fn(C.byref(cfg), n, *[x.data_ptr() for x in ins], *[o.data_ptr() for o in outs], stream), name)
    torch.cuda.synchronize()                                         # the no-alloc form does not synchronise itself
    return [o.cpu().numpy() for o in outs]


@gpu
def test_noalloc_reach_verify_matches_allocating_entry_point():
    import torch

    from human_robot_gym_b200 import cabi
    from human_robot_gym_b200.config import ShieldConfig
    L = cabi.load_library()
    torch.cuda.set_device(0)
    # two calls in a row on the shared array: a large Schunk batch, then a smaller Panda batch (other J, other layout)
    for robot, n, seed in (("schunk", 4000, 501), ("panda", 1237, 502), ("schunk", 33, 503)):
        files = pu.default_files(robot)
        sc = ShieldConfig.from_files(0.004, *files)
        cfg = cabi.make_cfg(sc, 8)
        J, H = cfg.nb_joints, cfg.n_ext + 2 * cfg.n_bodies
        ins = [torch.as_tensor(np.ascontiguousarray(a), device="cuda:0") for a in _inputs(sc, files, n, seed)]
        want = _call(L.hrg_stage_reach_verify, "hrg_stage_reach_verify", cfg, n, ins, J, H)
        got = _call(L.hrg_stage_reach_verify_noalloc, "hrg_stage_reach_verify_noalloc", cfg, n, ins, J, H)
        for w, g, what in zip(want, got, ("robot capsules", "human capsules", "min_dist", "is_safe")):
            assert not np.isnan(w.astype(np.float64)).any(), what
            if what == "is_safe":
                assert np.array_equal(w, g), what
            else:
                assert np.abs(w - g).max() < 1e-12, what            # the tolerance of test_parity_gaps (north_star: 1e-5)
        assert 0.05 < want[3].mean() < 0.95                          # both verdicts occur


@gpu
def test_noalloc_rejects_batches_beyond_the_static_array():
    import torch

    from human_robot_gym_b200 import cabi
    from human_robot_gym_b200.config import ShieldConfig
    L = cabi.load_library()
    torch.cuda.set_device(0)
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    cfg = cabi.make_cfg(sc, 8)
    n = 6042                                                         # include/hrg_shield.h: at most 6 041 Schunk envs
    z = torch.zeros(n * 130, dtype=torch.float64, device="cuda:0")
    u = torch.zeros(n, dtype=torch.uint8, device="cuda:0")
    rc = L.hrg_stage_reach_verify_noalloc(C.byref(cfg), n, *([z.data_ptr()] * 8), z.data_ptr(), z.data_ptr(), z.data_ptr(),
                                          u.data_ptr(), None)
    assert rc == -1                                                  # HRG_EINVAL
    msg = L.hrg_last_error().decode()
    assert "6041" in msg and "hrg_stage_reach_verify_noalloc" in msg
    rc = L.hrg_stage_reach_verify_noalloc(C.byref(cfg), 6041, *([z.data_ptr()] * 8), None, None, None, u.data_ptr(), None)
    torch.cuda.synchronize()
    assert rc == 0
