"""Regenerates tests/golden/plot_capsules.npz by IMPORTING the reference's own PlotCapsule (run in the build container).

HRG/controllers/failsafe_controller/failsafe_controller/plot_capsule.py is plain Python (math, numpy, scipy): it is loaded
from /root/reference by file path and evaluated on seeded capsules - random ones, spheres (p1 == p2), nearly degenerate
axes, axes along +z / -z (the shortest-arc quaternion's singular direction) - so the committed outputs are the reference's.
"""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("HRG_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "human_robot_gym/controllers/failsafe_controller/failsafe_controller/plot_capsule.py")


def main():
    spec = importlib.util.spec_from_file_location("ref_plot_capsule", SRC)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    rng = np.random.default_rng(2024)
    n = 240
    caps = np.zeros((n, 7))
    caps[:, :3] = rng.uniform(-2, 2, (n, 3))
    caps[:, 3:6] = caps[:, :3] + rng.uniform(-1, 1, (n, 3))
    caps[:, 6] = rng.uniform(0.01, 0.4, n)
    caps[200:210, 3:6] = caps[200:210, :3]                                   # spheres
    caps[210:220, 3:6] = caps[210:220, :3] + rng.uniform(-4e-7, 4e-7, (10, 3))  # |axis| around the 1e-6 switch
    caps[220:225, 3:6] = caps[220:225, :3] + np.array([0, 0, 1.0]) * rng.uniform(0.1, 1, (5, 1))   # +z
    caps[225:230, 3:6] = caps[225:230, :3] - np.array([0, 0, 1.0]) * rng.uniform(0.1, 1, (5, 1))   # -z: norm_q = 0
    caps[230:240, 3:6] = caps[230:240, :3] + np.array([1e-9, -1e-9, -1.0]) * rng.uniform(0.1, 1, (10, 1))  # almost -z
    pos, size, mat = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros((n, 3, 3))
    for i in range(n):
        pc = mod.PlotCapsule(caps[i, 0:3].tolist(), caps[i, 3:6].tolist(), float(caps[i, 6]))
        pos[i], size[i], mat[i] = pc.pos, pc.size, pc.mat
    np.savez_compressed(os.path.join(HERE, "plot_capsules.npz"), caps=caps, pos=pos, size=size, mat=mat)
    print("wrote plot_capsules.npz", n, "capsules")


if __name__ == "__main__":
    sys.exit(main())
