"""The operation counter behind profiles/r03_flop_count.md (oracle/flopcount: basic-block counts of a
`-fsanitize-coverage=trace-pc` build joined with objdump) must count known loops exactly, and the committed counts of the
reference must be the ones bench.py quotes. CPU only; needs gcc + objdump, not /root/reference."""
import ctypes as C
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r"""
#include <math.h>
double dot(const double* a, const double* b, int n) { double s = 0; for (int i = 0; i < n; ++i) s += a[i] * b[i]; return s; }
double norm3(double x, double y, double z) { return sqrt(x * x + y * y + z * z); }
double horner(const double* c, int n, double x) { double p = c[0]; for (int i = 1; i <= n; ++i) p = p * x + c[i]; return p; }
double ratio(double a, double b) { return a > b ? a / b : fmin(a, b) - b; }
"""


@pytest.mark.skipif(not (shutil.which("gcc") and shutil.which("objdump")), reason="needs gcc and objdump")
def test_block_counter_counts_known_loops_exactly(tmp_path):
    from oracle.flopcount import count_flops as cf
    (tmp_path / "t.c").write_text(SRC)
    run = lambda *a: subprocess.run(a, check=True, cwd=tmp_path)
    run("gcc", "-O2", "-fPIC", "-c", os.path.join(ROOT, "oracle", "flopcount", "cov_rt.c"), "-o", "cov_rt.o")   # NOT instrumented
    run("gcc", "-O2", "-fPIC", "-fsanitize-coverage=trace-pc", "-c", "t.c", "-o", "t.o")
    run("gcc", "-shared", "-o", "libt.so", "t.o", "cov_rt.o", "-lm", "-ldl")
    lib = str(tmp_path / "libt.so")
    L = cf.bind(C.CDLL(lib))
    d = C.c_double
    L.dot.restype = L.norm3.restype = L.horner.restype = L.ratio.restype = d
    L.dot.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.norm3.argtypes = [d, d, d]
    L.horner.argtypes = [C.c_void_p, C.c_int, d]
    L.ratio.argtypes = [d, d]
    a = np.arange(1000.0)
    assert L.dot(a.ctypes.data, a.ctypes.data, 1000) == float((a * a).sum())       # disabled: nothing is counted
    L.cov_enable(1)
    L.cov_phase(0)
    L.dot(a.ctypes.data, a.ctypes.data, 1000)
    L.cov_phase(1)
    for _ in range(10):
        L.norm3(1.0, 2.0, 3.0)
    L.cov_phase(2)
    L.horner(a.ctypes.data, 6, 0.5)
    L.cov_phase(3)
    for _ in range(7):
        L.ratio(3.0, 2.0)           # the division branch only
    L.cov_enable(0)
    L.dot(a.ctypes.data, a.ctypes.data, 1000)
    blocks, entry = cf.parse_disassembly(lib)
    assert {"dot", "norm3", "horner", "ratio"} <= set(entry)
    flop = lambda t: {k: v for k, v in t.items() if k in cf.FLOP_CLASSES}
    assert flop(cf.phase_totals(L, blocks, 0)) == {"mul": 1000, "add": 1000}
    assert flop(cf.phase_totals(L, blocks, 1)) == {"mul": 30, "add": 20, "sqrt": 10}
    assert flop(cf.phase_totals(L, blocks, 2)) == {"mul": 6, "add": 6}
    t3 = cf.phase_totals(L, blocks, 3)
    assert flop(t3) == {"div": 7} and t3["cmp"] == 7                                 # the untaken branch costs nothing
    L.cov_reset()
    assert not cf.phase_totals(L, blocks, 0)


def test_committed_reference_counts_are_consistent():
    """profiles/r03_flop_count_*.json: totals add up, the FK cross-check holds, bench.py reads the same figure."""
    import bench
    got = {}
    for tag in ("schunk_ssm", "panda_ssm", "schunk_pfl", "schunk_ssm_config4", "schunk_pfl_config4"):
        with open(os.path.join(ROOT, "profiles", f"r03_flop_count_{tag}.json")) as f:
            r = json.load(f)
        p = r["per_step"]
        assert abs(r["F_step"] - sum(p.get(k, 0.0) for k in ("add", "mul", "div", "sqrt", "minmax"))) < 1e-6
        assert abs(sum(v["flop"] for v in r["per_phase"].values()) - r["F_step"]) < 1.0     # driver code is the difference
        assert abs(sum(v["flop"] for g, v in r["per_group"].items() if not g.startswith("driver")) - r["F_step"]) < 1e-6
        assert r["calls_per_step"]["SafetyShield::step"] == 1.0 and r["calls_per_step"]["LongTermTraj::interpolate"] == 3.0
        assert sum(r["branch_hist"]) == r["shield_steps"]
        got[tag] = r["F_step"]
    with open(os.path.join(ROOT, "profiles", "r03_flop_count_schunk_ssm.json")) as f:
        r = json.load(f)
    # FK of the reference: 2 motions x 6 joints x 2 products of 4x4 matrices (112 operations) and as many matrix-vector ones (28)
    assert r["per_group"]["Eigen shim: dense matrix / vector products"]["flop"] == 24 * 112 + 24 * 28
    assert 9.0e3 < got["schunk_ssm"] < 1.0e4 < got["panda_ssm"] < got["schunk_pfl"] < got["schunk_pfl_config4"]
    args = bench.parse_args([])
    fc = bench.reference_flop_count(args)
    assert fc is not None and fc["F_step"] == got["schunk_ssm"]
    assert bench.reference_flop_count(bench.parse_args(["--robot", "panda"]))["F_step"] == got["panda_ssm"]
    assert bench.reference_flop_count(bench.parse_args(["--robot", "nosuchrobot"])) is None
