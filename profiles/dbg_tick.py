import sys, faulthandler, time, os
sys.path.insert(0, os.getcwd())
faulthandler.dump_traceback_later(70, exit=True)
import numpy as np, torch
from tests import parity_util as pu
from human_robot_gym_b200.batched import BatchedSafetyShield
from human_robot_gym_b200.config import ShieldConfig
from human_robot_gym_b200.synthetic import SyntheticScenario
robot=sys.argv[1]; n=int(sys.argv[2])
files=pu.default_files(robot)
sc=ShieldConfig.from_files(0.004,*files)
scen=SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 1002)
g=BatchedSafetyShield(n,0.004,*files,init_qpos=scen.init_q,device="cuda:0",ltt_capacity=1024)
phase=np.arange(n)%25
for k in range(80):
    t=k*0.004
    g.humanMeasurement(torch.from_numpy(scen.keypoints(t)),t)
    gm=phase==(k%25)
    g.newLongTermTrajectory(torch.from_numpy(scen.next_goal(gm)),env_mask=torch.from_numpy(gm))
    g.step(t)
    torch.cuda.synchronize()
    print("tick",k,"ok", g.stats()["n_replan"], flush=True)
print("done")
