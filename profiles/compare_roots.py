"""Derivative-chain root finder (csrc/ltp_roots_fast.cuh) against the exact restatement of the reference's QR iteration on
harvested polynomials: classification (finite / none) and value. usage: python profiles/compare_roots.py polys.npz [...]"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import numpy as np  # noqa: E402

from tests.host_emul import emul  # noqa: E402

emul._LIB = os.environ.get("HRG_EMUL_LIB", "/tmp/emul_tr/libhrg_emul.so")
L = emul.lib()
L.emul_poly_root_fast_batch.argtypes = [C.c_longlong, C.c_int] + [C.c_void_p] * 4
MARGIN = float(os.environ.get("MARGIN", 1e4))
tot = {}
for f in sys.argv[1:]:
    P = np.load(f)["polys"]
    for nc in (5, 6, 7):
        c = np.ascontiguousarray(P[P[:, 0] == nc][:, 1:1 + nc])
        if not len(c):
            continue
        c = np.unique(c, axis=0)
        n = len(c)
        fast, marg, reg = np.zeros(n), np.zeros(n), np.zeros(n)
        t0 = time.perf_counter()
        L.emul_poly_root_fast_batch(n, nc, c.ctypes.data, fast.ctypes.data, marg.ctypes.data, None)
        t1 = time.perf_counter()
        L.emul_poly_root_fast_batch(n, nc, c.ctypes.data, fast.ctypes.data, marg.ctypes.data, reg.ctypes.data)
        t2 = time.perf_counter()
        amb = marg < MARGIN
        fin_f, fin_r = np.isfinite(fast), np.isfinite(reg)
        cls_bad = (fin_f != fin_r) & ~amb
        both = fin_f & fin_r & ~amb
        rel = np.abs(fast[both] - reg[both]) / np.abs(reg[both])
        val_bad = both.copy()
        val_bad[both] = rel > 1e-9
        print(f"{os.path.basename(f)} deg {nc - 1}: {n} distinct, ambiguous {amb.sum()} ({amb.mean():.2e}), class mismatch {cls_bad.sum()}, "
              f"value rel diff max {rel.max() if len(rel) else 0:.2e} p99.9 {np.percentile(rel, 99.9) if len(rel) else 0:.2e}, >1e-9: {val_bad.sum()}; "
              f"fast {1e9 * (t1 - t0) / n:.0f} ns/poly, QR {1e9 * (t2 - t1 - (t1 - t0)) / n:.0f} ns/poly")
        # what the margin of the disagreements looks like (all of them, also the ambiguous ones)
        dis = (fin_f != fin_r) | (fin_f & fin_r & (np.abs(fast - reg) > 1e-9 * np.abs(reg)))
        if dis.any():
            m = marg[dis]
            print("   disagreements incl. ambiguous:", int(dis.sum()), "margins: max %.3g" % m.max(), "sorted top:", np.sort(m)[-5:])
            for i in np.flatnonzero(dis & ~amb)[:5]:
                print("   !!", c[i].tolist(), "fast", fast[i], "qr", reg[i], "margin", marg[i])
