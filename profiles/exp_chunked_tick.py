"""Experiment: one tick of 65 536 envs as C independent env blocks, each on its own stream (replan chains overlap)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root
import numpy as np, torch
import bench
from human_robot_gym_b200.batched import BatchedSafetyShield
from human_robot_gym_b200.config import ShieldConfig
from tests.parity_util import default_files, mocap_joint_names

class A: pass
args = A(); args.ring = 32; args.seed = 1003
files = default_files("schunk"); sc = ShieldConfig.from_files(0.004, *files); names = mocap_joint_names(files[2])
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
dev = torch.device("cuda", 0)
scen, kp_h, goals_h, masks_h = bench.make_inputs(args, sc, N, 1003, names)
kp_d = torch.from_numpy(kp_h).to(dev); goals_d = torch.from_numpy(goals_h).to(dev); masks_d = torch.from_numpy(masks_h).to(dev)
W, S = 50, 100
total = W + S + 8
t_d = (torch.arange(total, dtype=torch.float64, device=dev) * 0.004)[:, None].expand(total, N).contiguous()
for C in [int(c) for c in (sys.argv[2].split(",") if len(sys.argv) > 2 else "1,2,4,8".split(","))]:
    bounds = [(i * N // C, (i + 1) * N // C) for i in range(C)]
    parts = [BatchedSafetyShield(hi - lo, 0.004, *files, init_qpos=scen.init_q[lo:hi], device=dev, ltt_capacity=512, config=sc) for lo, hi in bounds]
    streams = [torch.cuda.Stream(device=dev) for _ in bounds]
    # contiguous per-part views of the inputs
    kp_p = [kp_d[:, lo:hi].contiguous() for lo, hi in bounds]
    g_p = [goals_d[:, lo:hi].contiguous() for lo, hi in bounds]
    m_p = [masks_d[:, lo:hi].contiguous() for lo, hi in bounds]
    t_p = [t_d[:, lo:hi].contiguous() for lo, hi in bounds]
    main = torch.cuda.current_stream(dev)
    ev_start = torch.cuda.Event(); ev_done = [torch.cuda.Event() for _ in bounds]
    def tick(k):
        ev_start.record(main)
        for i, (p, st) in enumerate(zip(parts, streams)):
            st.wait_event(ev_start)
            with torch.cuda.stream(st):
                p.tick_device(kp_p[i][bench.ring_index(k, 32)], t_p[i][k], m_p[i][k % 25], g_p[i][(k // 25) % g_p[i].shape[0]])
                ev_done[i].record(st)
        for e in ev_done:
            main.wait_event(e)
    for k in range(W): tick(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(W, W + S): tick(k)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / S
    chk = sum(float(p.q.sum().item()) for p in parts)
    print(json.dumps({"envs": N, "chunks": C, "ms_per_tick": ms, "steps_per_s": N / ms * 1e3, "checksum": chk}), flush=True)
    for p in parts: p.close()
    del parts
