"""The same operation count as oracle/flopcount/count_flops.py (profiles/r03_flop_count.md), taken on the DEVICE's math:
csrc/*.cuh compiled for the host by tests/host_emul (env by env, sequential planner ladder) with
`-fsanitize-coverage=trace-pc`, driven by the bench workload. Says where the device's formulation of the tick spends more
or less arithmetic than the reference's compiled code. Measurement tooling; nothing here is on a product path.
  python profiles/flop_count_device_math.py [--robot schunk] [--shield ssm] [--envs 2048] [--ticks 100] [--out x.json]"""
import argparse
import collections
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from oracle.flopcount import count_flops as cf  # noqa: E402
from tests.host_emul import emul  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--robot", default="schunk")
ap.add_argument("--mocap", default="mujoco_mocap")
ap.add_argument("--traj", default=None)
ap.add_argument("--shield", default="ssm")
ap.add_argument("--envs", type=int, default=2048)
ap.add_argument("--warmup", type=int, default=25)
ap.add_argument("--ticks", type=int, default=100)
ap.add_argument("--root-mode", type=int, default=1)
ap.add_argument("--out", default=None)
a = ap.parse_args()

d = tempfile.mkdtemp(prefix="emul_cov_")
he = os.path.join(ROOT, "tests", "host_emul")
subprocess.check_call(["gcc", "-O2", "-fPIC", "-c", os.path.join(ROOT, "oracle", "flopcount", "cov_rt.c"), "-o", d + "/cov_rt.o"])
subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-w", "-fsanitize-coverage=trace-pc",
                       "-o", d + "/libhrg_emul.so", os.path.join(he, "emul.cc"), d + "/cov_rt.o", "-lm", "-ldl"], cwd=he)
emul._LIB = d + "/libhrg_emul.so"
os.utime(emul._LIB)
L = cf.bind(emul.lib())

import bench  # noqa: E402
from human_robot_gym_b200 import workloads  # noqa: E402
from human_robot_gym_b200.config import ShieldConfig  # noqa: E402
from tests import parity_util as pu  # noqa: E402

bargs = bench.parse_args(["--envs", str(a.envs), "--robot", a.robot, "--mocap", a.mocap, "--shield", a.shield]
                         + (["--traj", a.traj] if a.traj else []))
files = workloads.default_files(bargs.robot, bargs.mocap, bargs.traj)
sc = ShieldConfig.from_files(bench.SAMPLE_TIME, *files)
names = workloads.mocap_joint_names(files[2])
n = a.envs
scen, kp, goals, masks = bench.make_inputs(bargs, sc, n, bargs.seed + 7919, names)
emul.set_root_mode(a.root_mode)
sh = pu.EmulShield(n, bench.SAMPLE_TIME, *files, base_pose=np.zeros((n, 6)), init_qpos=scen.init_q,
                   shield_type=2 if a.shield == "pfl" else 1, ltt_capacity=bargs.ltt_capacity)
for k in range(a.warmup + a.ticks):
    L.cov_enable(int(k >= a.warmup))
    L.cov_phase(0)
    sh.humanMeasurement(kp[bench.ring_index(k, bargs.ring)], np.full(n, k * bench.SAMPLE_TIME))
    L.cov_phase(1)
    sh.newLongTermTrajectory(goals[(k // bench.GOAL_PERIOD) % goals.shape[0]], env_mask=masks[k % bench.GOAL_PERIOD])
    L.cov_phase(2)
    sh.step(np.full(n, k * bench.SAMPLE_TIME))
L.cov_enable(0)

blocks, entry = cf.parse_disassembly(emul._LIB)
dem = cf.demangle(sorted({b["fn"] for b in blocks.values()}))
steps = n * a.ticks
tot, per_fn, calls = collections.Counter(), collections.defaultdict(collections.Counter), collections.Counter()
per_phase = {}
for ph, name in enumerate(cf.PHASES):
    pt = collections.Counter()
    for pc, c in cf.read_phase(L, ph):
        b = blocks.get(int(pc))
        if b is None:
            continue
        c = int(c)
        f = dem[b["fn"]].split("(")[0]
        for k2, v in b["ops"].items():
            pt[k2] += v * c
            per_fn[f][k2] += v * c
        for k2, v in b["libm"].items():
            pt["libm:" + k2] += v * c
        if entry.get(b["fn"]) == int(pc):
            calls[f] += c
    per_phase[name] = {k2: v / steps for k2, v in sorted(pt.items())}
    tot.update(pt)
flop = lambda cn: sum(cn.get(k2, 0) for k2 in cf.FLOP_CLASSES)
res = {
    "what": "FP64 operations per shield-step of the DEVICE's math (csrc/*.cuh compiled for the host, tests/host_emul, "
            f"root_mode {a.root_mode}), same counting method and workload as r03_flop_count_*.json",
    "workload": f"{a.robot} {a.shield.upper()} {a.mocap}, {n} envs x {a.ticks} ticks after {a.warmup}",
    "replans_per_step": sh.stats()["n_replan"] / (n * (a.warmup + a.ticks)),
    "F_step": flop(tot) / steps,
    "per_step": {k2: v / steps for k2, v in sorted(tot.items())},
    "per_phase": {p: dict(v, flop=sum(v.get(k2, 0) for k2 in cf.FLOP_CLASSES)) for p, v in per_phase.items()},
    "functions_flop_per_step": [[f, flop(cn) / steps, calls.get(f, 0) / steps]
                                for f, cn in sorted(per_fn.items(), key=lambda kv: -flop(kv[1]))[:40] if flop(cn) > 0],
}
js = json.dumps(res, indent=1)
print(js)
if a.out:
    open(a.out, "w").write(js + "\n")
