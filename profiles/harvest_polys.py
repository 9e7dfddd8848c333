"""Harvest every polynomial the planner solves on the bench workload (host emulation built with -DHRG_POLY_LOG
-DHRG_TRACE_PREFILL -DHRG_HOST_USE_RR into /tmp/emul_tr) and compare the derivative-chain root finder with the exact QR one.
  python profiles/harvest_polys.py [envs] [ticks] [out.npz]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.getcwd())
import numpy as np  # noqa: E402

from tests.host_emul import emul  # noqa: E402

emul._LIB = os.environ.get("HRG_EMUL_LIB", "/tmp/emul_tr/libhrg_emul.so")
os.utime(emul._LIB)
import bench  # noqa: E402
from human_robot_gym_b200 import workloads  # noqa: E402
from human_robot_gym_b200.config import ShieldConfig  # noqa: E402
from tests import parity_util as pu  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 110
out = sys.argv[3] if len(sys.argv) > 3 else "/tmp/emul_tr/polys.npz"
robot = os.environ.get("ROBOT", "schunk")
args = bench.parse_args(["--envs", str(n), "--robot", robot])
files = workloads.default_files(robot)
sc = ShieldConfig.from_files(0.004, *files)
names = workloads.mocap_joint_names(files[2])
scen, kp, goals, masks = bench.make_inputs(args, sc, n, int(os.environ.get("SEED", 1003)), names)
sh = pu.EmulShield(n, 0.004, *files, base_pose=np.zeros((n, 6)), init_qpos=scen.init_q, ltt_capacity=512)
L = emul.lib()
L.emul_poly_log.restype = C.c_longlong
L.emul_poly_log.argtypes = [C.c_int, C.c_void_p, C.c_longlong]
L.emul_poly_log(1, None, 0)
for k in range(ticks):
    sh.humanMeasurement(kp[bench.ring_index(k, args.ring)], np.full(n, k * 0.004))
    sh.newLongTermTrajectory(goals[(k // 25) % goals.shape[0]], env_mask=masks[k % 25])
    sh.step(np.full(n, k * 0.004))
rows = L.emul_poly_log(1, None, 0)
buf = np.zeros((rows, 8))
L.emul_poly_log(0, buf.ctypes.data, rows)
print("polynomials logged:", rows, "replans", sh._st["n_replan"])
np.savez_compressed(out, polys=buf)
