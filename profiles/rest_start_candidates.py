"""Host experiment: the timeScaling candidates whose root is a close call (margin < kRootMargin) on the bench workload -
how many are the exact-rest structure x^k q(x), what the QR iteration returns for them, and whether
ltp_cand_rejected_from_rest (csrc/ltp_scaling.cuh) ever contradicts the outcome of the exact path.
  g++ ... -DHRG_HOST_USE_RR -DHRG_CAND_LOG -DHRG_RULE_DEBUG -include cstdio -o /tmp/emul_tr/libhrg_emul_cl.so tests/host_emul/emul.cc
  python profiles/rest_start_candidates.py <envs> <ticks>"""
import sys, os, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np
from tests.host_emul import emul
emul._LIB = '/tmp/emul_tr/libhrg_emul_cl.so'
os.utime(emul._LIB)
import bench
from human_robot_gym_b200 import workloads
from human_robot_gym_b200.config import ShieldConfig
from tests import parity_util as pu
n = int(sys.argv[1]); ticks = int(sys.argv[2])
args = bench.parse_args(["--envs", str(n)])
files = workloads.default_files("schunk")
sc = ShieldConfig.from_files(0.004, *files)
names = workloads.mocap_joint_names(files[2])
scen, kp, goals, masks = bench.make_inputs(args, sc, n, 1003, names)
sh = pu.EmulShield(n, 0.004, *files, base_pose=np.zeros((n,6)), init_qpos=scen.init_q, ltt_capacity=512)
L = emul.lib()
for k in range(ticks):
    sh.humanMeasurement(kp[bench.ring_index(k, args.ring)], np.full(n, k*0.004))
    sh.newLongTermTrajectory(goals[(k // 25) % goals.shape[0]], env_mask=masks[k % 25])
    sh.step(np.full(n, k*0.004))
out = (C.c_longlong*16)()
L.emul_cand_log(out)
names = ["cand polys","ambiguous","amb: >=2 trailing zeros","amb: other","  of trailing: rest start","  deflated deg>2","  QR gave garbage","  true root accepted","  garbage rep accepted","  QR outcome accepted","  predictable reject","  !! predictable but QR accepted","  garbage AND accepted","  garbage > 1e-2","  rule says rejected","  !! rule says rejected but QR accepted"]
for nm, v in zip(names, out): print(f"{nm:36s} {v}")
print("replans", sh._st)
