"""Shared helpers of the parity tests, smoke() and bench.py's checker legs (test infrastructure)."""
from __future__ import annotations

import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from human_robot_gym_b200.workloads import (config_dir, configs, default_files, full_body_mocap,  # noqa: E402,F401
                                            mocap_joint_names)


def variant_files(robot="schunk", mocap=None, traj=None):
    """default_files(robot) with keys of the mocap and / or trajectory YAML overridden (written to a fresh temporary
    directory): configurations the shipped files never exercise, e.g. a measurement delay and measurement errors
    (all shipped mocap files carry zeros), per-joint limits that differ from joint to joint, other PFL parameters."""
    import tempfile

    import yaml
    t, r, m = default_files(robot)
    d = tempfile.mkdtemp(prefix="shield_cfg_variant_")
    if mocap:
        cfg = dict(configs()["mujoco_mocap"], **mocap)
        m = os.path.join(d, "mujoco_mocap_variant.yaml")
        with open(m, "w") as f:
            yaml.safe_dump(cfg, f, default_flow_style=None)
    if traj:
        cfg = dict(configs()[f"trajectory_parameters_{robot}"], **traj)
        t = os.path.join(d, f"trajectory_parameters_{robot}_variant.yaml")
        with open(t, "w") as f:
            yaml.safe_dump(cfg, f, default_flow_style=None)
    return t, r, m


def make_checker(files, n, base_pose, init_q, sample_time=0.004, shield_type=1, threads=8):
    """The CPU checker for a batch: the reference build (oracle/_ref). There is no second checker."""
    from oracle import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref/libsara_ref.so is not built: run `python __graft_entry__.py` (or `make -C oracle "
                           "ref`) in a container that has /root/reference; the prebuilt library travels to the GPU box")
    return ref.RefBatch(n, sample_time, files[0], files[1], files[2], base_pose, init_q, shield_type=shield_type,
                        threads=threads), "reference"


class AdaptivePulseScenario:
    """A scenario that provokes branch 1 of determineNextMotion (SS/src/safety_shield.cc:521-527: safe, but the recovery
    path lags the failsafe path). That branch needs a SAFE verdict on a tick where the recovery-path planning failed
    (recovery_path_correct_ false with a stale recovery path behind the failsafe path), which only happens for a few ticks
    in the middle of a failsafe braking manoeuvre. So the human stands ON the robot (unsafe: the robot brakes), and steps
    6 m away exactly when the REFERENCE's recovery path has gone stale while the failsafe path still moves; after a pause
    the human comes back. The release time is read from the reference's state (`chk.paths()`), and the resulting keypoints
    are fed identically to both sides - the CUDA path is never consulted."""

    def __init__(self, n_envs, joint_names, q_min, q_max, seed, goal_period=60, goal_step=0.8):
        from human_robot_gym_b200.synthetic import rest_pose
        rng = np.random.default_rng(seed)
        self.n, self.rng = int(n_envs), rng
        self.q_min, self.q_max = np.asarray(q_min, float), np.asarray(q_max, float)
        self.J = self.q_min.shape[0]
        base = rest_pose(joint_names)
        self.near = base[None] + np.array([0.05, 0.0, -0.4]) + rng.uniform(-0.05, 0.05, (self.n, 1, 3))
        self.far = base[None] + np.array([6.0, 0.0, 0.0]) + np.zeros((self.n, 1, 3))
        self.init_q = np.zeros((self.n, self.J))
        self._goal = self.init_q.copy()
        self.goal_period, self.goal_step = goal_period, goal_step
        self.state = np.zeros(self.n, int)             # 0 = away (timer running), 1 = on the robot, waiting for the trigger
        self.timer = 20 + rng.integers(0, 30, self.n)
        self.prev_rec = None

    def goal_mask(self, k):
        return np.full(self.n, k % self.goal_period == 0)

    def next_goal(self, mask=None):
        step = self.rng.uniform(-self.goal_step, self.goal_step, (self.n, self.J))
        self._goal = np.clip(self._goal + step, self.q_min, self.q_max)
        return self._goal.copy()

    def keypoints_adaptive(self, k, chk):
        P = chk.paths()
        rec, fs_vel = P[:, 0, 6:9], P[:, 1, 7]
        trig = np.zeros(self.n, bool) if self.prev_rec is None else (
            (np.abs(rec - self.prev_rec).max(axis=1) == 0) & (fs_vel > 0.05))
        self.prev_rec = rec.copy()
        rel = (self.state == 1) & trig
        self.state[rel] = 0
        self.timer[rel] = 40 + self.rng.integers(0, 20, int(rel.sum()))
        self.timer[self.state == 0] -= 1
        self.state[(self.state == 0) & (self.timer <= 0)] = 1
        return np.where((self.state == 1)[:, None, None], self.near, self.far)


def random_base_pose(rng, n):
    """(x, y, z in +-1 m, roll / pitch / yaw in +-pi) per env - the six init_* arguments of the reference constructor
    (SS/src/robot_reach.cc:5-17) that FailsafeController fills from the robot's base (failsafe_controller.py:160-189)."""
    return np.concatenate([rng.uniform(-1.0, 1.0, (n, 3)), rng.uniform(-np.pi, np.pi, (n, 3))], axis=1)


class EmulShield:
    """BatchedSafetyShield's verbs over tests/host_emul (the SAME csrc/*.cuh tick logic compiled for the host, env by env):
    lets the GPU-less container run closed_loop_compare's scenarios (base poses, mid-episode resets, mixed shield types,
    branch 1) against the reference before any GPU time is spent. Test infrastructure; never loaded by the product."""

    def __init__(self, n_envs, sample_time, traj_file, robot_file, mocap_file, base_pose=None, init_qpos=None,
                 shield_type=1, ltt_capacity=1024, **_):
        from human_robot_gym_b200.cabi import make_cfg
        from human_robot_gym_b200.config import ShieldConfig
        from tests.host_emul import emul
        sc = ShieldConfig.from_files(sample_time, traj_file, robot_file, mocap_file)
        self.em = emul.EmulBatch(make_cfg(sc, ltt_capacity), n_envs, base_pose, init_qpos, shield_type)
        self.H, self.n_envs = self.em.H, n_envs
        self._st = {"n_steps": 0, "n_replan": 0, "n_ltt_truncated": 0, "n_pfl_rejected": 0}

    def reset(self, base_pose, init_qpos, current_time=0.0, shield_type=1, env_mask=None):
        self.em.reset(base_pose, init_qpos, shield_type, mask=env_mask)

    def humanMeasurement(self, kp, t, env_mask=None):
        self.em.human_measurement(kp, np.broadcast_to(np.asarray(t, float), (self.n_envs,)), mask=env_mask)

    def newLongTermTrajectory(self, g, env_mask=None):
        self.em.new_goal(g, mask=env_mask)

    def step(self, t):
        o = self.em.step(t)
        for k, v in o.items():
            setattr(self, k, v)
        self._st["n_steps"] += self.n_envs
        self._st["n_replan"] += int(((o["status"] & 1) != 0).sum())
        self._st["n_ltt_truncated"] += int(((o["status"] & 16) != 0).sum())

    def capsules(self):
        return self.em.capsules()

    def stats(self):
        return dict(self._st)

    def close(self):
        self.em = None


def closed_loop_compare(files, n_envs=256, ticks=100, seed=1001, device="cuda:0", sample_time=0.004, shield_type=1,
                        goal_period=25, stagger=True, ltt_capacity=1024, base_pose=None, collect=None,
                        store_capsules=True, human=None, scenario=None, reset_period=0, reset_types=None,
                        backend="cuda", meas_every=1, meas_lag=0.0):
    """Drives BatchedSafetyShield (CUDA) and the CPU checker in lockstep on identical synthetic inputs.

    Decisions (is_safe, branch) must match exactly; an env whose FIRST mismatch happens on a tick where either side
    reports a capsule distance within 1e-6 m of 0 is counted as near-threshold and excluded (north_star). Envs are
    dropped from the comparison after their first mismatch (their closed loops have diverged).

    store_capsules=False is the product configuration: reach + verify then runs as the model cascade (a human model that
    is not needed for the verdict is not evaluated, its min_dist stays +inf), and only the distances of the models that
    were evaluated are compared.

    shield_type: one ShieldType for all envs or an [n] array (mixed OFF / SSM / PFL batches).
    base_pose: None (zeros), "random" (random_base_pose per env), or an array broadcastable to [n, 6].
    human: None for the synthetic human, else an object with checker_keypoints(k) -> [n,K,3] numpy (the checker's input)
    and device_keypoints(k) -> [n,K,3] CUDA tensor (the CUDA path's input), e.g. the recorded-animation replay.
    scenario: None (SyntheticScenario) or a class with the same constructor, e.g. AdaptivePulseScenario.
    backend: "cuda" = BatchedSafetyShield through the C ABI (the parity tests proper); "host_emul" = EmulShield (CPU tests).
    reset_period > 0: every reset_period ticks a random third of the envs is reset MID-EPISODE on both sides
    (SafetyShield::reset, SS/src/safety_shield.cc:164-201) with a new random base pose, a new init_qpos and - when
    reset_types is a list of ShieldTypes - a new random shield type out of it; the human keypoint history is kept, as
    HumanReach::reset keeps it (SS/src/human_reach.cc:44-50).
    meas_every > 1: on every tick each env receives a human measurement with probability 1 / meas_every only (masked
    humanMeasurement; the reach time t_command - last_meas_timestep_ + t_brake of SS/src/human_reach.cc:76-80 then differs
    from env to env); meas_lag: the measurement's timestamp lies that far behind the cycle time handed to step().
    """
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario

    if backend == "cuda":
        import torch

        from human_robot_gym_b200.batched import BatchedSafetyShield as Shield
        to_dev = lambda x: torch.from_numpy(np.ascontiguousarray(x))            # noqa: E731
        to_np = lambda x: None if x is None else x.cpu().numpy()                # noqa: E731
    else:
        Shield = EmulShield
        to_dev = to_np = lambda x: x                                             # noqa: E731
    sc = ShieldConfig.from_files(sample_time, *files)
    Scen = scenario or SyntheticScenario
    scen = Scen(n_envs, mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, seed)
    rng = np.random.default_rng(seed + 99991)
    rng_meas = np.random.default_rng(seed + 7)
    if base_pose is None:
        pose = np.zeros((n_envs, 6))
    elif isinstance(base_pose, str) and base_pose == "random":
        pose = random_base_pose(rng, n_envs)
    else:
        pose = np.broadcast_to(np.asarray(base_pose, float), (n_envs, 6))
    pose = np.ascontiguousarray(pose)
    types = np.ascontiguousarray(np.broadcast_to(np.asarray(shield_type, np.int32), (n_envs,))).copy()
    chk, kind = make_checker(files, n_envs, pose, scen.init_q, sample_time, types)
    gpu = Shield(n_envs, sample_time, *files, base_pose=pose, init_qpos=scen.init_q,
                 shield_type=types, device=device, ltt_capacity=ltt_capacity, store_capsules=store_capsules)
    H = gpu.H
    alive = np.ones(n_envs, bool)
    n_mis, n_near, max_sp, max_md, max_caps = 0, 0, 0.0, 0.0, 0.0
    hist = np.zeros(3, np.int64)
    unsafe, n_reset, n_verified = 0, 0, 0
    phase = (np.arange(n_envs) % goal_period) if stagger else np.zeros(n_envs, int)
    q_lo, q_hi = np.asarray(sc.traj.q_min_allowed, float), np.asarray(sc.traj.q_max_allowed, float)
    for k in range(ticks):
        t = k * sample_time
        if reset_period and k > 0 and k % reset_period == 0:
            rm = rng.random(n_envs) < 1.0 / 3.0
            new_pose = np.where(rm[:, None], random_base_pose(rng, n_envs), pose)
            new_q = np.clip(rng.uniform(-1.0, 1.0, (n_envs, q_lo.shape[0])), q_lo, q_hi)
            if reset_types is not None:
                types = np.where(rm, rng.choice(np.asarray(reset_types, np.int32), n_envs), types).astype(np.int32)
            pose = np.ascontiguousarray(new_pose)
            chk.reset(pose, new_q, np.full(n_envs, t), types, mask=rm)
            gpu.reset(pose, new_q, t, types, env_mask=to_dev(rm))
            if hasattr(scen, "_goal"):
                scen._goal[rm] = new_q[rm]   # the goal random walk restarts from the new configuration
            n_reset += int(rm.sum())
        if human is not None:
            kp = human.checker_keypoints(k)
        elif hasattr(scen, "keypoints_adaptive"):
            kp = scen.keypoints_adaptive(k, chk)
        else:
            kp = scen.keypoints(t - meas_lag)
        mm = None
        if meas_every > 1:
            mm = rng_meas.random(n_envs) < 1.0 / meas_every
            if k == 0:
                mm[:] = True
        if mm is None or mm.any():
            chk.human_measurement(kp, t - meas_lag, mask=mm)
            gpu.humanMeasurement(to_dev(kp) if human is None else human.device_keypoints(k), t - meas_lag,
                                 env_mask=None if mm is None else to_dev(mm))
        gmask = scen.goal_mask(k) if hasattr(scen, "goal_mask") else (phase == (k % goal_period))
        if gmask.any():
            g = scen.next_goal(gmask)
            chk.new_goal(g, mask=gmask)
            gpu.newLongTermTrajectory(to_dev(g), env_mask=to_dev(gmask))
        a = chk.step(t)
        gpu.step(t)
        b = {key: to_np(getattr(gpu, key)) for key in ("q", "dq", "ddq", "dddq", "is_safe", "branch", "status")}
        r_g, h_g, md_g = [to_np(x) for x in gpu.capsules()]
        r_c, h_c, md_c = chk.capsules(H)
        bad = ((a["is_safe"] != b["is_safe"]) | (a["branch"] != b["branch"])) & alive
        if bad.any():
            near = ((b["status"] & 32) != 0) | (np.abs(np.where(np.isfinite(md_c), md_c, 1.0)).min(axis=1) < 1e-6)
            n_near += int((bad & near).sum())
            n_mis += int((bad & ~near).sum())
            alive &= ~bad
        if alive.any():
            for key in ("q", "dq", "ddq", "dddq"):
                max_sp = max(max_sp, float(np.abs(a[key][alive] - b[key][alive]).max()))
            # the reference keeps the capsules of its last VERIFIED step (stale on joint-limit / OFF steps): compare the
            # envs whose step ran reach + verify on the CUDA side (finite distances) and is not OFF
            verified = alive & (types != 0)
            if store_capsules:
                ver = verified & np.isfinite(md_g).all(axis=1) & np.isfinite(md_c).all(axis=1)
                if ver.any():
                    n_verified += int(ver.sum())
                    max_md = max(max_md, float(np.abs(md_g[ver] - md_c[ver]).max()))
                    max_caps = max(max_caps, float(np.abs(r_g[ver] - r_c[ver]).max()),
                                   float(np.abs(h_g[ver] - h_c[ver]).max()))
            else:
                both = verified[:, None] & np.isfinite(md_g) & np.isfinite(md_c)
                if both.any():
                    n_verified += int(both.any(axis=1).sum())
                    max_md = max(max_md, float(np.abs(md_g[both] - md_c[both]).max()))
            hist += np.bincount(a["branch"][alive], minlength=3)[:3]
            unsafe += int((a["is_safe"][alive] == 0).sum())
        if collect is not None:
            collect(k, a, b)
    stats = gpu.stats()
    gpu.close()
    return {"checker": kind, "n_envs": n_envs, "ticks": ticks, "mismatch_excl_near": n_mis, "near_threshold_excluded": n_near,
            "max_setpoint_diff": max_sp, "max_min_dist_diff": max_md, "max_capsule_diff": max_caps,
            "branch_hist": hist.tolist(), "unsafe_frac": unsafe / max(1, int(hist.sum())), "n_reset": n_reset,
            "n_verified": n_verified, "n_alive": int(alive.sum()), "gpu_stats": stats}


def assert_parity(res, n_envs=None):
    """The north_star bar on a closed_loop_compare result: decisions exact (near-threshold cases reported and bounded),
    setpoints / distances / capsules within 1e-5, no truncated trajectory, no rejected PFL reset."""
    n = n_envs or res["n_envs"]
    assert res["mismatch_excl_near"] == 0, res
    assert res["max_setpoint_diff"] < 1e-5 and res["max_min_dist_diff"] < 1e-5 and res["max_capsule_diff"] < 1e-5, res
    assert res["near_threshold_excluded"] <= max(1, n // 100), res
    assert res["gpu_stats"]["n_ltt_truncated"] == 0 and res["gpu_stats"]["n_pfl_rejected"] == 0, res
    assert res["gpu_stats"]["n_steps"] == n * res["ticks"], res
