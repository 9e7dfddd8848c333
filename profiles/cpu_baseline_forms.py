"""The CPU baseline forms BASELINE.md section 4 lists beside the all-cores figure of bench.py:
  (a) the reference's per-env loop on ONE host thread (4 096 envs x 100 ticks of the bench workload), steps/s;
  (c) BASELINE config 1 - ONE env stepped through the Python boundary in the call pattern of human_env.py:503-524
      (humanMeasurement, newLongTermTrajectory every 25th tick, step): per-tick latency percentiles of replan / other ticks,
      for the compiled reference (oracle/_ref through ctypes) and for the drop-in safety_shield_py on the GPU.
Test / measurement infrastructure (imports oracle/): python profiles/cpu_baseline_forms.py > profiles/r03_cpu_baseline_forms.json"""
import json
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import numpy as np  # noqa: E402

import bench  # noqa: E402
from human_robot_gym_b200 import workloads  # noqa: E402
from human_robot_gym_b200.config import ShieldConfig  # noqa: E402
from human_robot_gym_b200.synthetic import SyntheticScenario  # noqa: E402
from oracle import ref  # noqa: E402

files = workloads.default_files("schunk")
sc = ShieldConfig.from_files(0.004, *files)
names = workloads.mocap_joint_names(files[2])
out = {}
args = bench.parse_args(["--cpu-envs", "4096"])
cb, ms, done = bench.cpu_loop(args, files, sc, names, 4096, 25, 100, 1, budget_s=60.0)
out["single_thread"] = {"value": cb["value"], "unit": "shield-steps/sec", "cores": 1, "sample": cb["sample"]}


def pct(x):
    x = np.asarray(x) * 1e6
    return {"n": len(x), "p50_us": float(np.percentile(x, 50)), "p90_us": float(np.percentile(x, 90)),
            "p99_us": float(np.percentile(x, 99)), "max_us": float(x.max())}


def one_env(make, T=3000):
    scen = SyntheticScenario(1, names, sc.traj.q_min_allowed, sc.traj.q_max_allowed, 7)
    sh, meas, goal, step = make(scen)
    rep, oth = [], []
    for k in range(T):
        t = k * 0.004
        kp = scen.keypoints(t)[0]
        t0 = time.perf_counter()
        meas(sh, kp, t)
        if k % 25 == 0:
            goal(sh, scen.next_goal()[0])
        step(sh, t)
        dt = time.perf_counter() - t0
        if k >= 50:
            (rep if k % 25 == 0 else oth).append(dt)
    return {"ticks_with_new_goal": pct(rep), "other_ticks": pct(oth)}


def make_ref(scen):
    sh = ref.RefBatch(1, 0.004, *files, np.zeros(6), scen.init_q)
    return (sh, lambda s, kp, t: s.human_measurement(kp[None], t), lambda s, g: s.new_goal(g[None]), lambda s, t: s.step(t))


out["config1_reference_through_ctypes"] = one_env(make_ref)
try:
    import torch
    if torch.cuda.is_available():
        from safety_shield_py import SafetyShield

        def make_gpu(scen):
            sh = SafetyShield(0.004, *files, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, list(scen.init_q[0]))
            return (sh, lambda s, kp, t: s.humanMeasurement([list(p) for p in kp], t),
                    lambda s, g: s.newLongTermTrajectory(list(g), [0.0] * len(g)), lambda s, t: s.step(t))
        out["config1_drop_in_safety_shield_py_on_gpu"] = one_env(make_gpu, T=1500)
except Exception as ex:  # reported, not hidden
    out["config1_drop_in_safety_shield_py_on_gpu"] = {"error": repr(ex)}
print(json.dumps(out, indent=1))
