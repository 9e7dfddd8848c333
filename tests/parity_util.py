"""Shared helpers of the parity tests, smoke() and bench.py's checker legs (test infrastructure)."""
from __future__ import annotations

import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from human_robot_gym_b200.workloads import (config_dir, configs, default_files, full_body_mocap,  # noqa: E402,F401
                                            mocap_joint_names)


def make_checker(files, n, base_pose, init_q, sample_time=0.004, shield_type=1, threads=8):
    """The CPU checker for a batch: the reference build (oracle/_ref). There is no second checker."""
    from oracle import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref/libsara_ref.so is not built: run `python __graft_entry__.py` (or `make -C oracle "
                           "ref`) in a container that has /root/reference; the prebuilt library travels to the GPU box")
    return ref.RefBatch(n, sample_time, files[0], files[1], files[2], base_pose, init_q, shield_type=shield_type,
                        threads=threads), "reference"


def closed_loop_compare(files, n_envs=256, ticks=100, seed=1001, device="cuda:0", sample_time=0.004, shield_type=1,
                        goal_period=25, stagger=True, ltt_capacity=1024, base_pose=None, collect=None,
                        store_capsules=True, human=None):
    """Drives BatchedSafetyShield (CUDA) and the CPU checker in lockstep on identical synthetic inputs.

    Decisions (is_safe, branch) must match exactly; an env whose FIRST mismatch happens on a tick where either side
    reports a capsule distance within 1e-6 m of 0 is counted as near-threshold and excluded (north_star). Envs are
    dropped from the comparison after their first mismatch (their closed loops have diverged).

    store_capsules=False is the product configuration: reach + verify then runs as the model cascade (a human model that
    is not needed for the verdict is not evaluated, its min_dist stays +inf), and only the distances of the models that
    were evaluated are compared.

    human: None for the synthetic human, else an object with checker_keypoints(k) -> [n,K,3] numpy (the checker's input)
    and device_keypoints(k) -> [n,K,3] CUDA tensor (the CUDA path's input), e.g. the recorded-animation replay.
    """
    import torch

    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario

    sc = ShieldConfig.from_files(sample_time, *files)
    scen = SyntheticScenario(n_envs, mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, seed)
    pose = np.zeros((n_envs, 6)) if base_pose is None else np.broadcast_to(np.asarray(base_pose, float), (n_envs, 6))
    pose = np.ascontiguousarray(pose)
    chk, kind = make_checker(files, n_envs, pose, scen.init_q, sample_time, shield_type)
    gpu = BatchedSafetyShield(n_envs, sample_time, *files, base_pose=pose, init_qpos=scen.init_q,
                              shield_type=shield_type, device=device, ltt_capacity=ltt_capacity, store_capsules=store_capsules)
    H = gpu.H
    alive = np.ones(n_envs, bool)
    n_mis, n_near, max_sp, max_md, max_caps = 0, 0, 0.0, 0.0, 0.0
    hist = np.zeros(3, np.int64)
    unsafe = 0
    phase = (np.arange(n_envs) % goal_period) if stagger else np.zeros(n_envs, int)
    for k in range(ticks):
        t = k * sample_time
        kp = scen.keypoints(t) if human is None else human.checker_keypoints(k)
        chk.human_measurement(kp, t)
        gpu.humanMeasurement(torch.from_numpy(kp) if human is None else human.device_keypoints(k), t)
        gmask = (phase == (k % goal_period))
        if gmask.any():
            g = scen.next_goal(gmask)
            chk.new_goal(g, mask=gmask)
            gpu.newLongTermTrajectory(torch.from_numpy(g), env_mask=torch.from_numpy(gmask))
        a = chk.step(t)
        gpu.step(t)
        b = {"q": gpu.q.cpu().numpy(), "dq": gpu.dq.cpu().numpy(), "ddq": gpu.ddq.cpu().numpy(),
             "dddq": gpu.dddq.cpu().numpy(), "is_safe": gpu.is_safe.cpu().numpy(), "branch": gpu.branch.cpu().numpy(),
             "status": gpu.status.cpu().numpy()}
        r_g, h_g, md_g = [None if x is None else x.cpu().numpy() for x in gpu.capsules()]
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
            if store_capsules:
                ver = alive & np.isfinite(md_g).all(axis=1) & np.isfinite(md_c).all(axis=1)
                if ver.any():
                    max_md = max(max_md, float(np.abs(md_g[ver] - md_c[ver]).max()))
                    max_caps = max(max_caps, float(np.abs(r_g[ver] - r_c[ver]).max()),
                                   float(np.abs(h_g[ver] - h_c[ver]).max()))
            else:
                both = alive[:, None] & np.isfinite(md_g) & np.isfinite(md_c)
                if both.any():
                    max_md = max(max_md, float(np.abs(md_g[both] - md_c[both]).max()))
            hist += np.bincount(a["branch"][alive], minlength=3)[:3]
            unsafe += int((a["is_safe"][alive] == 0).sum())
        if collect is not None:
            collect(k, a, b)
    stats = gpu.stats()
    gpu.close()
    return {"checker": kind, "n_envs": n_envs, "ticks": ticks, "mismatch_excl_near": n_mis, "near_threshold_excluded": n_near,
            "max_setpoint_diff": max_sp, "max_min_dist_diff": max_md, "max_capsule_diff": max_caps,
            "branch_hist": hist.tolist(), "unsafe_frac": unsafe / max(1, int(hist.sum())), "gpu_stats": stats}
