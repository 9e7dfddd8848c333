"""Debug aid: a short closed loop on the GPU with a watchdog (dumps the Python stack and exits if a kernel hangs).
  [HRG_DEBUG_SYNC=1] python profiles/dbg_tick.py <robot> <n_envs> [ticks]"""
import faulthandler
import os
import sys

sys.path.insert(0, os.getcwd())
faulthandler.dump_traceback_later(70, exit=True)
from tests import parity_util as pu  # noqa: E402

robot, n = sys.argv[1], int(sys.argv[2])
ticks = int(sys.argv[3]) if len(sys.argv) > 3 else 80
res = pu.closed_loop_compare(pu.default_files(robot), n_envs=n, ticks=ticks, seed=1002, store_capsules=False)
print(res)
pu.assert_parity(res)
print("done")
