"""CPU parity stress: the tick logic of csrc/*.cuh compiled for the host (tests/host_emul, root_mode 1) against the compiled
reference (oracle/_ref) in lockstep, 20 minutes over the nine robots the reference ships configs for, SSM / PFL, random base
poses, mid-episode resets on every fourth run. Measurement tooling (uses the checker); log: r03u_stress_parity_cpu.log."""
import sys, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests import parity_util as pu
t0=time.time(); tot=0; rep=0; worst=0.0; bad=[]
seed=9000
while time.time()-t0 < 1200:
    for robot in ("schunk","panda","iiwa","jaco","kinova3","sawyer","ur5e","modrob0","modrob1"):
        seed+=1
        st = [1,2,1][seed%3]
        bp = "random" if seed%2 else None
        res = pu.closed_loop_compare(pu.default_files(robot), n_envs=128, ticks=250, seed=seed, shield_type=st, base_pose=bp,
                                     store_capsules=False, backend="host_emul", ltt_capacity=4096,
                                     reset_period=60 if seed%4==0 else 0)
        tot += res["gpu_stats"]["n_steps"]; rep += res["gpu_stats"]["n_replan"]; worst=max(worst,res["max_setpoint_diff"])
        if res["mismatch_excl_near"] or res["near_threshold_excluded"] > 1 or res["max_setpoint_diff"] > 1e-7:
            bad.append((robot, seed, st, res["mismatch_excl_near"], res["near_threshold_excluded"], res["max_setpoint_diff"]))
            print("BAD", bad[-1], flush=True)
    print(f"{time.time()-t0:.0f}s steps {tot} replans {rep} worst {worst:.2e} bad {len(bad)}", flush=True)
print("DONE steps", tot, "replans", rep, "worst", worst, "bad", bad)
