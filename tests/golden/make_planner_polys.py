"""Writes tests/golden/planner_polys.npz: a sample of the polynomials LongTermPlanner hands to roots() on the bench workload.

1. host emulation with the polynomial log (every optSwitchTimes quartic incl. the ones k_replan_plan solves ahead of need,
   every timeScaling candidate of degree 4 / 5 / 6):
     g++ -O2 -std=c++17 -fPIC -shared -ffp-contract=off -w -DHRG_HOST_USE_RR -DHRG_TRACE_PREFILL -DHRG_POLY_LOG \
         -o /tmp/emul_tr/libhrg_emul.so tests/host_emul/emul.cc -lm
2. ROBOT=schunk SEED=1003 python profiles/harvest_polys.py 16384 150 /tmp/emul_tr/polys_schunk_1003.npz   (and SEED=77)
3. python tests/golden/make_planner_polys.py /tmp/emul_tr/polys_schunk_1003.npz /tmp/emul_tr/polys_schunk_77.npz
The fixture keeps 30 000 distinct quartics and 6 000 distinct polynomials of degree 5 and 6 each (seeded choice)."""
import os
import sys

import numpy as np

out = {}
rng = np.random.default_rng(0)
files = sorted(sys.argv[1:])
for deg in (4, 5, 6):
    parts = []
    for f in files:
        P = np.load(f)["polys"]
        parts.append(np.unique(P[P[:, 0] == deg + 1][:, 1:deg + 2], axis=0))
    c = np.unique(np.concatenate(parts), axis=0)
    keep = {4: 30000, 5: 6000, 6: 6000}[deg]
    out[f"deg{deg}"] = c[rng.permutation(len(c))[:keep]]
    print(deg, len(c), "->", len(out[f"deg{deg}"]))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "planner_polys.npz"), **out)
