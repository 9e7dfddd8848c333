"""ncu target: W warm ticks of the bench workload, then C ticks between cudaProfilerStart/Stop.
  ncu --profile-from-start off --set full --import-source on --clock-control none -o out python profiles/ncu_tick.py [envs] [warm] [capture]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.getcwd())
import bench  # noqa: E402
from human_robot_gym_b200 import workloads  # noqa: E402
from human_robot_gym_b200.batched import BatchedSafetyShield  # noqa: E402
from human_robot_gym_b200.config import ShieldConfig  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
W = int(sys.argv[2]) if len(sys.argv) > 2 else 60
C = int(sys.argv[3]) if len(sys.argv) > 3 else 1
shield_kind = sys.argv[4] if len(sys.argv) > 4 else "ssm"
args = bench.parse_args(["--envs", str(n), "--shield", shield_kind])
files = workloads.default_files("schunk")
sc = ShieldConfig.from_files(bench.SAMPLE_TIME, *files)
names = workloads.mocap_joint_names(files[2])
dev = torch.device("cuda:0")
scen, kp_h, goals_h, masks_h = bench.make_inputs(args, sc, n, 1003, names)
shield = BatchedSafetyShield(n, bench.SAMPLE_TIME, *files, init_qpos=scen.init_q, device=dev, ltt_capacity=512, config=sc,
                             shield_type=2 if shield_kind == "pfl" else 1)
kp_d, goals_d, masks_d = (torch.from_numpy(x).to(dev) for x in (kp_h, goals_h, masks_h))
total = W + C + 4
t_d = (torch.arange(total, dtype=torch.float64, device=dev) * bench.SAMPLE_TIME)[:, None].expand(total, n).contiguous()


def tick(k):
    shield.tick_device(kp_d[bench.ring_index(k, args.ring)], t_d[k], masks_d[k % bench.GOAL_PERIOD],
                       goals_d[(k // bench.GOAL_PERIOD) % goals_d.shape[0]])


for k in range(W):
    tick(k)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for k in range(W, W + C):
    tick(k)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("stats", shield.stats(clear=True))
