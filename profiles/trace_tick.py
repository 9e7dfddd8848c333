"""Per-task timeline of the replan chain of ONE steady-state tick (profiling build, profiles/build_trace.sh).
  python profiles/trace_tick.py [envs] [warm] [out.npz]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.getcwd())
from human_robot_gym_b200 import cabi  # noqa: E402

cabi.LIB_PATH = os.path.join(os.getcwd(), "profiles", "_trace", "libhrg_shield_trace.so")
import bench  # noqa: E402
from human_robot_gym_b200 import workloads  # noqa: E402
from human_robot_gym_b200.batched import BatchedSafetyShield  # noqa: E402
from human_robot_gym_b200.config import ShieldConfig  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
W = int(sys.argv[2]) if len(sys.argv) > 2 else 60
out = sys.argv[3] if len(sys.argv) > 3 else None
args = bench.parse_args(["--envs", str(n)])
files = workloads.default_files("schunk")
sc = ShieldConfig.from_files(bench.SAMPLE_TIME, *files)
names = workloads.mocap_joint_names(files[2])
dev = torch.device("cuda:0")
scen, kp_h, goals_h, masks_h = bench.make_inputs(args, sc, n, 1003, names)
shield = BatchedSafetyShield(n, bench.SAMPLE_TIME, *files, init_qpos=scen.init_q, device=dev, ltt_capacity=512, config=sc)
kp_d, goals_d, masks_d = (torch.from_numpy(x).to(dev) for x in (kp_h, goals_h, masks_h))
total = W + 8
t_d = (torch.arange(total, dtype=torch.float64, device=dev) * bench.SAMPLE_TIME)[:, None].expand(total, n).contiguous()
L = cabi.load_library()
L.hrg_trace_read.argtypes = [C.c_void_p]
L.hrg_trace_read_spans.argtypes = [C.c_void_p]
buf = np.zeros((4, 1 << 15, 4), dtype=np.uint64)
spans = np.zeros((16, 2), dtype=np.uint64)
SPAN_NAMES = {0: "k_replan_plan", 1: "k_replan_deep", 2: "k_replan_sample", 4: "k_measure", 5: "k_pre", 6: "k_plan", 7: "k_reach_verify p0",
              8: "k_verify_stage<1>", 9: "k_verify_stage<2>", 10: "k_select (not replanned)", 11: "k_reach_verify p1", 12: "k_select (last)"}


def tick(k):
    shield.tick_device(kp_d[bench.ring_index(k, args.ring)], t_d[k], masks_d[k % bench.GOAL_PERIOD],
                       goals_d[(k // bench.GOAL_PERIOD) % goals_d.shape[0]])


for k in range(W):
    tick(k)
torch.cuda.synchronize()
res = []
for k in range(W, W + 3):
    assert L.hrg_trace_read(buf.ctypes.data) == 0 and L.hrg_trace_read_spans(spans.ctypes.data) == 0
    tick(k)
    torch.cuda.synchronize()
    assert L.hrg_trace_read(buf.ctypes.data) == 0 and L.hrg_trace_read_spans(spans.ctypes.data) == 0
    sp = spans.astype(np.int64)
    ok = sp[:, 1] > 0
    z0 = sp[ok, 0].min()
    print(f"tick {k} timeline (us from the first block of the tick):")
    for i in np.argsort(np.where(ok, sp[:, 0], np.iinfo(np.int64).max)):
        if ok[i]:
            print(f"    {SPAN_NAMES.get(int(i), i):26s} {(sp[i, 0] - z0) / 1e3:7.1f} -> {(sp[i, 1] - z0) / 1e3:7.1f}  ({(sp[i, 1] - sp[i, 0]) / 1e3:6.1f})")
    res.append(buf.copy())
    t00 = None
    r3 = buf[3]
    m3 = r3[:, 3] > 0
    if m3.any():
        print(f"tick {k} k_replan_plan phases (us, mean / max): checkInputs + quartic roots {r3[m3, 0].mean() / 1e3:.1f} / {r3[m3, 0].max() / 1e3:.1f}  "
              f"optSwitchTimes {r3[m3, 1].mean() / 1e3:.1f} / {r3[m3, 1].max() / 1e3:.1f}  closed-form candidates {r3[m3, 2].mean() / 1e3:.1f} / {r3[m3, 2].max() / 1e3:.1f}")
    for ki, name in enumerate(("plan", "deep", "sample+tail")):
        r = buf[ki]
        m = r[:, 1] > 0
        if not m.any():
            continue
        t0, t1 = r[m, 0].astype(np.int64), r[m, 1].astype(np.int64)
        if t00 is None:
            t00 = t0.min()
        d = (t1 - t0) / 1e3
        q = np.percentile(d, [50, 90, 99, 100])
        sm = r[m, 2].astype(np.int64)
        per_sm = np.bincount(sm, minlength=148)
        print(f"tick {k} {name:11s} tasks {m.sum():5d} start {(t0.min() - t00) / 1e3:7.1f} us  last start {(t0.max() - t00) / 1e3:7.1f}  "
              f"end {(t1.max() - t00) / 1e3:7.1f}  task us mean {d.mean():6.1f} p50 {q[0]:6.1f} p90 {q[1]:6.1f} p99 {q[2]:6.1f} max {q[3]:6.1f}"
              f"  tasks/SM min {per_sm.min()} max {per_sm.max()}  extra max {int(r[m, 3].max())}")
        # the slowest tasks
        order = np.argsort(-d)[:5]
        print("      slowest:", [(int(np.flatnonzero(m)[i]), round(float(d[i]), 1), int(sm[i]), int(r[m, 3][i])) for i in order])
cnt = np.zeros(4, dtype=np.uint64)
L.hrg_trace_read_counts.argtypes = [C.c_void_p]
L.hrg_trace_read_counts(cnt.ctypes.data)
st = shield.stats()
print(f"exact QR solves over {W + 3} ticks ({st['n_replan']} replans): deep phase B {cnt[0]}, inline close calls {cnt[1]}; "
      f"candidates settled by the rest-start rule {cnt[2]}")
if out:
    np.savez_compressed(out, trace=np.stack(res))
