"""Large closed-loop parity runs of the CUDA path (root_mode 1: derivative-chain roots, rest-start rule) against the compiled
reference, beyond what the test suite holds: python profiles/stress_parity.py > profiles/r03_stress_parity.log"""
import os
import sys
import time

sys.path.insert(0, os.getcwd())
from tests import parity_util as pu  # noqa: E402

THREADS = max(1, min(32, os.cpu_count() or 1))
chk0 = pu.make_checker
pu.make_checker = lambda *a, **k: chk0(*a, **dict(k, threads=THREADS))
total = 0
for robot, n, ticks, seed, kw in (("schunk", 65536, 150, 9001, {}), ("panda", 32768, 150, 9002, {}),
                                  ("schunk", 32768, 200, 9003, dict(base_pose="random", reset_period=40)),
                                  ("schunk", 16384, 150, 9004, dict(shield_type=2))):
    t0 = time.time()
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=n, ticks=ticks, seed=seed, store_capsules=False, ltt_capacity=512, **kw)
    pu.assert_parity(res)
    total += res["gpu_stats"]["n_replan"]
    print(f"{robot} {n} envs x {ticks} ticks seed {seed} {kw}: mismatches excl. near {res['mismatch_excl_near']}, near-threshold excluded "
          f"{res['near_threshold_excluded']}, max setpoint diff {res['max_setpoint_diff']:.3g}, max min_dist diff {res['max_min_dist_diff']:.3g}, "
          f"replans {res['gpu_stats']['n_replan']} (failed {res['gpu_stats']['n_replan_failed']}), branches {res['branch_hist']}, "
          f"alive {res['n_alive']}/{n}, {time.time() - t0:.0f} s", flush=True)
print("total replans compared:", total)
