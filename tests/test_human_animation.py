"""Animation -> keypoints: `_control_human` + the kinematic tree of human.xml, on the device.

The checker is oracle/human_fk_oracle.py (MuJoCo's mj_kinematics restated with quaternions; parity unpinned - MuJoCo is
absent, see the oracle's header); the device code uses the rotation-matrix form of the same tree, so the two derivations
check each other. Inputs are the committed fixtures tests/golden/human_tree.json and cmu_clips.npz (the head of two of
the reference's recorded CMU clips). Tolerance: 1e-5 m (north_star); what is observed is < 1e-12.
"""
import json
import math
import os

import numpy as np
import pytest

from oracle import human_fk_oracle as hfo
from tests.host_emul import emul

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-5  # metres, north_star


def load_fixture():
    from human_robot_gym_b200.human_animation import load_clips_npz, load_tree_json
    return (load_tree_json(os.path.join(HERE, "golden", "human_tree.json")),
            load_clips_npz(os.path.join(HERE, "golden", "cmu_clips.npz")))


def random_placements(rng, n, center=(0.0, 0.0, 0.0), spread=1.0):
    pos = np.asarray(center) + rng.uniform(-spread, spread, (n, 3)) * np.array([1.0, 1.0, 0.0])
    yaw = rng.uniform(0, 2 * np.pi, n)
    quat = np.stack([np.cos(yaw / 2), np.zeros(n), np.zeros(n), np.sin(yaw / 2)], axis=1)  # (w,x,y,z)
    return pos, quat


def oracle_frame(tree, clips, clip, frame, pos_off, rot_off):
    anim, info = clips[clip]
    p, q = hfo.root_pose(anim, info, frame, pos_off, rot_off)
    angles = {k: v[frame] for k, v in anim.items() if not k.startswith("Pelvis")}
    return hfo.site_positions(tree, p, q, angles)


def test_fixture_is_the_reference_model():
    tree, clips = load_fixture()
    from human_robot_gym_b200.human_animation import JOINT_ELEMENTS
    assert tree["names"] == JOINT_ELEMENTS and len(tree["names"]) == 23
    assert tree["parent"][:3] == [-1, -1, -1] and all(p < i for i, p in enumerate(tree["parent"]))
    assert clips[0][0]["Pelvis_quat"].shape == (400, 4) and clips[1][0]["L_Hip_x"].shape == (60,)
    assert len(clips[0][0]) == 73  # 3 + 1 + 69 channels, as in the reference's pickles


def test_rest_pose_and_bone_lengths():
    """Zero angles: sites = root frame applied to the anchors. Any frame: a site keeps its distance to the parent's site
    (the bone is rigid) - a property of the tree that holds whatever MuJoCo's rounding is."""
    tree, clips = load_fixture()
    c = np.asarray(tree["anchor"])
    zero = {n + a: 0.0 for n in tree["names"] for a in ("_x", "_y", "_z")}
    p0, q0 = np.array([0.3, -0.2, 0.9]), np.array([math.cos(0.4), 0.0, 0.0, math.sin(0.4)])
    s = hfo.site_positions(tree, p0, q0, zero)
    assert np.abs(s - (p0 + c @ hfo._qmat(q0).T)).max() < 1e-14
    for fr in (0, 57, 399):
        s = oracle_frame(tree, clips, 0, fr, [0, 0, 0], [1, 0, 0, 0])
        # the root-pose chain of `_control_human` puts the recorded (y-up) human upright on the floor of the z-up world
        names = tree["names"]
        assert -0.05 < s[names.index("L_Toe")][2] < 0.2 and 1.3 < s[names.index("Head")][2] < 1.6
        for b, p in enumerate(tree["parent"]):
            if p >= 0:
                assert abs(np.linalg.norm(s[b] - s[p]) - np.linalg.norm(c[b] - c[p])) < 1e-12


def test_device_arithmetic_on_host_matches_oracle():
    """kernels_human.cuh compiled for the host (tests/host_emul) against the oracle on recorded frames."""
    tree, clips = load_fixture()
    from human_robot_gym_b200.human_animation import pack_clip
    rng = np.random.default_rng(7)
    n = 200
    pos, quat = random_placements(rng, n)
    which = rng.integers(0, 2, n)
    frame = np.array([rng.integers(0, clips[c][0]["L_Hip_x"].shape[0]) for c in which])
    banks = [pack_clip(a, tree["names"]) for a, _ in clips]
    rows = np.stack([banks[c][f] for c, f in zip(which, frame)])
    info = np.stack([clips[c][1]["position_offset"] + clips[c][1]["orientation_quat"] for c in which]).astype(float)
    got = emul.human_keypoints(tree, rows, info, np.concatenate([pos, quat], axis=1))
    want = np.stack([oracle_frame(tree, clips, c, f, pos[i], quat[i]) for i, (c, f) in enumerate(zip(which, frame))])
    assert np.abs(got - want).max() < 1e-12
    # random (large) angles exercise every hinge, not only what the recording happens to contain
    rows2 = rows.copy()
    rows2[:, 7:] = rng.uniform(-np.pi, np.pi, rows2[:, 7:].shape)
    rows2[:, 3:7] = rng.normal(size=(n, 4))  # unnormalised Pelvis quaternions: both sides normalise
    got2 = emul.human_keypoints(tree, rows2, info, np.concatenate([pos, quat], axis=1))
    for i in range(0, n, 10):
        anim = {"Pelvis_pos_x": rows2[i:i + 1, 0], "Pelvis_pos_y": rows2[i:i + 1, 1], "Pelvis_pos_z": rows2[i:i + 1, 2],
                "Pelvis_quat": rows2[i:i + 1, 3:7]}
        p, q = hfo.root_pose(anim, clips[which[i]][1], 0, pos[i], quat[i])
        ang = {nm + "_" + a: rows2[i, 7 + 3 * b + j] for b, nm in enumerate(tree["names"]) for j, a in enumerate("xyz")}
        assert np.abs(got2[i] - hfo.site_positions(tree, p, q, ang)).max() < 1e-12


def test_tree_validation():
    tree, _ = load_fixture()
    bad = dict(tree, parent=[1] + tree["parent"][1:])  # child before parent
    with pytest.raises(ValueError):
        emul.human_keypoints(bad, np.zeros((1, 76)), np.array([[0, 0, 0, 0, 0, 0, 1.0]]), np.array([[0, 0, 0, 1.0, 0, 0, 0]]))


def test_replay_state_machine():
    """HumanReplay walks clip 1 (60 frames) then clip 0, at floor(t / step_length) like `_control_human`."""
    tree, clips = load_fixture()
    step = 500 / 120.0
    rp = hfo.HumanReplay(tree, clips, [1, 0], [0, 0, 0], [1, 0, 0, 0], step)
    seen = []
    for t in range(0, 300):
        rp.control_human(t)
        seen.append((rp.clip_ids[rp.index], rp.animation_time))
    assert seen[0] == (1, 0) and seen[5] == (1, 1)
    first0 = next(i for i, s in enumerate(seen) if s[0] == 0)
    assert seen[first0 - 1] == (1, 59) and seen[first0] == (0, 0) and math.floor(first0 / step) == 60


# ---------------------------------------------------------------------------------------------------------------- GPU
class _ReplayPair:
    """The recorded human for closed_loop_compare: oracle replays for the checker, the device replay for the CUDA path."""

    def __init__(self, tree, clips, n, seed, device, center, spread, list_len=2):
        import torch
        from human_robot_gym_b200.human_animation import BatchedHumanAnimation
        rng = np.random.default_rng(seed)
        self.n, self.torch = n, torch
        self.pos, self.quat = random_placements(rng, n, center, spread)
        self.ids = rng.integers(0, len(clips), (n, list_len)).astype(np.int32)
        self.cpu = [hfo.HumanReplay(tree, clips, self.ids[e], self.pos[e], self.quat[e], 500 / 120.0) for e in range(n)]
        self.gpu = BatchedHumanAnimation(tree, clips, n, device=device, n_animations_per_env=list_len)
        self.gpu.reset(self.pos, self.quat, self.ids)
        self.max_diff = 0.0
        self._last = None

    def checker_keypoints(self, k):
        self._last = np.stack([r.control_human(k) for r in self.cpu])
        return self._last

    def device_keypoints(self, k):
        kp = self.gpu.human_measurement(k)
        self.max_diff = max(self.max_diff, float(np.abs(kp.cpu().numpy() - self._last).max()))
        return kp


@pytest.mark.gpu
def test_device_keypoints_match_oracle_over_clip_changes():
    import torch
    tree, clips = load_fixture()
    n = 109  # ragged: 3 full tiles of 32 envs + 13
    pair = _ReplayPair(tree, clips, n, seed=11, device="cuda:0", center=(-1.5, 0.0, 0.0), spread=1.5)
    clip_changes = 0
    for k in list(range(0, 12)) + list(range(243, 258)) + list(range(1662, 1672)) + list(range(1912, 1924)):
        want = pair.checker_keypoints(k)
        got = pair.device_keypoints(k).cpu().numpy()
        assert np.abs(got - want).max() < TOL
        clip = pair.gpu.clip.cpu().numpy()
        frame = pair.gpu.frame.cpu().numpy()
        assert (clip == np.array([r.clip_ids[r.index] for r in pair.cpu])).all()
        assert (frame == np.array([r.animation_time for r in pair.cpu])).all()
        clip_changes += int((np.array([r.index for r in pair.cpu]) > 0).any())
    assert clip_changes > 0 and pair.max_diff < 1e-11
    # masked reset: only the masked envs start over
    mask = (np.arange(n) % 3 == 0)
    pair.gpu.reset(pair.pos, pair.quat, pair.ids, env_mask=mask)
    pair.gpu.human_measurement(0)
    fr = pair.gpu.frame.cpu().numpy()
    assert (fr[mask] == 0).all()
    assert pair.gpu.launch_count > 0
    pair.gpu.close()
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_recorded_human_closed_loop_matches_checker():
    """BASELINE config 1 widened to a batch: Schunk shields next to replayed CMU humans; device FK -> device shield against
    oracle FK -> reference shield. Decisions exact, setpoints < 1e-5."""
    from tests.parity_util import closed_loop_compare, default_files
    tree, clips = load_fixture()
    n = 64
    pair = _ReplayPair(tree, clips, n, seed=12, device="cuda:0", center=(-1.5, 0.0, 0.0), spread=0.6)
    res = closed_loop_compare(default_files("schunk"), n_envs=n, ticks=200, seed=1001, device="cuda:0", human=pair)
    assert res["mismatch_excl_near"] == 0, res
    assert res["max_setpoint_diff"] < TOL and res["max_min_dist_diff"] < TOL and pair.max_diff < 1e-11, res
    assert res["branch_hist"][0] > 0 and res["branch_hist"][2] > 0, res  # the human does come close
    pair.gpu.close()


@pytest.mark.gpu
def test_tick_replay_host_equals_keypoints_then_tick():
    """hrg_shield_tick_replay_host (host clock + goals in, setpoints out) = hrg_human_keypoints + the three shield calls."""
    import torch
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.human_animation import BatchedHumanAnimation
    from human_robot_gym_b200.synthetic import SyntheticScenario
    from tests.parity_util import default_files, mocap_joint_names
    tree, clips = load_fixture()
    files = default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n = 77
    scen = SyntheticScenario(n, mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 31)
    rng = np.random.default_rng(31)
    pos, quat = random_placements(rng, n, (-1.5, 0.0, 0.0), 0.7)
    ids = rng.integers(0, 2, (n, 2)).astype(np.int32)
    sh = [BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0") for _ in range(2)]
    hu = [BatchedHumanAnimation(tree, clips, n, device="cuda:0", n_animations_per_env=2) for _ in range(2)]
    for h in hu:
        h.reset(pos, quat, ids)
    pin = dict(pin_memory=True)
    clock_p, t_p = torch.zeros(n, dtype=torch.int64, **pin), torch.zeros(n, dtype=torch.float64, **pin)
    q_p, dq_p, ddq_p = (torch.zeros(n, 6, dtype=torch.float64, **pin) for _ in range(3))
    fl_p = torch.zeros(n, dtype=torch.uint8, **pin)
    for k in range(70):
        t = k * 0.004
        gmask = (np.arange(n) % 25) == (k % 25)
        g = scen.next_goal(gmask)
        kp = hu[0].human_measurement(k + 230)  # the clip change of the 60-frame clip falls into the run
        sh[0].humanMeasurement(kp, t)
        if gmask.any():
            sh[0].newLongTermTrajectory(torch.from_numpy(g), env_mask=torch.from_numpy(gmask))
        sh[0].step(t)
        clock_p.fill_(k + 230)
        t_p.fill_(t)
        gm_p, gq_p = torch.from_numpy(gmask.astype(np.uint8)).pin_memory(), torch.from_numpy(g).pin_memory()
        hu[1].tick_replay_host(sh[1], clock_p, t_p, gm_p if gmask.any() else None, gq_p if gmask.any() else None,
                               q_p, dq_p, ddq_p, fl_p)
        torch.cuda.synchronize()
        assert np.array_equal(q_p.numpy(), sh[0].q.cpu().numpy()) and np.array_equal(ddq_p.numpy(), sh[0].ddq.cpu().numpy())
        assert np.array_equal(fl_p.numpy() & 1, sh[0].is_safe.cpu().numpy())
    assert int((fl_p.numpy() >> 1).max()) <= 2
    for x in sh + hu:
        x.close()


@pytest.mark.gpu
def test_tick_replay_piped_equals_tick_replay_host():
    """hrg_shield_tick_replay_piped (setpoints of tick k copied on a second stream while tick k + 1 computes, two output slots)
    returns the setpoints of hrg_shield_tick_replay_host, tick for tick, when read one tick late."""
    import torch
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.human_animation import BatchedHumanAnimation
    from human_robot_gym_b200.synthetic import SyntheticScenario
    from tests.parity_util import default_files, mocap_joint_names
    tree, clips = load_fixture()
    files = default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n = 211
    scen = SyntheticScenario(n, mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 57)
    rng = np.random.default_rng(57)
    pos, quat = random_placements(rng, n, (-1.5, 0.0, 0.0), 0.7)
    ids = rng.integers(0, 2, (n, 2)).astype(np.int32)
    sh = [BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0") for _ in range(2)]
    hu = [BatchedHumanAnimation(tree, clips, n, device="cuda:0", n_animations_per_env=2) for _ in range(2)]
    for h in hu:
        h.reset(pos, quat, ids)
    pin = dict(pin_memory=True)
    mk = lambda: [torch.zeros(n, 6, dtype=torch.float64, **pin) for _ in range(3)] + [torch.zeros(n, dtype=torch.uint8, **pin)]  # noqa: E731
    ref_out, slots = mk(), [mk(), mk()]
    clock = [torch.zeros(n, dtype=torch.int64, **pin) for _ in range(3)]
    tt = [torch.zeros(n, dtype=torch.float64, **pin) for _ in range(3)]
    copy_stream = torch.cuda.Stream(device="cuda:0")
    want, evs = {}, [None, None]
    T = 60
    for k in range(T):
        gmask = (np.arange(n) % 25) == (k % 25)
        g = scen.next_goal(gmask)
        gm_p, gq_p = torch.from_numpy(gmask.astype(np.uint8)).pin_memory(), torch.from_numpy(g).pin_memory()
        clock[2].fill_(k + 230)
        tt[2].fill_(k * 0.004)
        hu[0].tick_replay_host(sh[0], clock[2], tt[2], gm_p, gq_p, *ref_out)
        torch.cuda.synchronize()
        want[k] = [x.clone() for x in ref_out]
        sl = k & 1
        clock[sl].fill_(k + 230)
        tt[sl].fill_(k * 0.004)
        hu[1].tick_replay_piped(sh[1], sl, clock[sl], tt[sl], gm_p, gq_p, *slots[sl], copy_stream)
        ev = torch.cuda.Event()
        ev.record(copy_stream)
        evs[sl] = ev
        if k >= 1:                                   # the setpoints of the PREVIOUS tick have arrived
            evs[sl ^ 1].synchronize()
            for a, b in zip(slots[sl ^ 1], want[k - 1]):
                assert torch.equal(a, b), k
        torch.cuda.synchronize()                     # (gm_p / gq_p are per-iteration buffers)
    for a, b in zip(slots[(T - 1) & 1], want[T - 1]):
        assert torch.equal(a, b)
    for x in sh + hu:
        x.close()
