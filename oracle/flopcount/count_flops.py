"""FP64 operations per shield-step of the REFERENCE, counted (TEST INFRASTRUCTURE; not on any product path).

SURVEY.md §8(d) and BASELINE.md §3 carry F_step as an estimate "to be replaced by an instrumented count of the oracle".
This is that count. `make -C oracle flopcount` compiles the unmodified reference a second time with
`-fsanitize-coverage=trace-pc` (a call at the head of every basic block; oracle/flopcount/cov_rt.c counts executions per
call site). This script drives the bench workload (bench.make_inputs: same scenario, same staggered goals) through that
library on one thread, then joins the block counts with `objdump -d` of the same file: every FP64 arithmetic instruction of
a block counts once per execution of the block. Calls into libm (sin, cos, pow, ...) are listed by name, their inside is
not counted. The instrumented code keeps the reference's arithmetic; the calls only stop gcc from vectorising loops, so
every operation shows up as a scalar instruction.

  python oracle/flopcount/count_flops.py [--robot schunk] [--mocap mujoco_mocap] [--shield ssm] [--envs 2048]
                                         [--warmup 25] [--ticks 100] [--out profiles/r03_flop_count.json]
"""
from __future__ import annotations

import argparse
import collections
import ctypes as C
import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

COV_LIB = os.path.join(ROOT, "oracle", "_ref", "libsara_ref_cov.so")
PHASES = ["humanMeasurement", "newLongTermTrajectory", "step"]

# x86-64 baseline (no AVX): what gcc emits for `double` arithmetic
CLASSES = [
    ("add", re.compile(r"^(addsd|subsd)$"), 1),
    ("mul", re.compile(r"^mulsd$"), 1),
    ("div", re.compile(r"^divsd$"), 1),
    ("sqrt", re.compile(r"^sqrtsd$"), 1),
    ("minmax", re.compile(r"^(minsd|maxsd)$"), 1),
    ("add", re.compile(r"^(addpd|subpd|haddpd|addsubpd)$"), 2),
    ("mul", re.compile(r"^mulpd$"), 2),
    ("div", re.compile(r"^divpd$"), 2),
    ("sqrt", re.compile(r"^sqrtpd$"), 2),
    ("minmax", re.compile(r"^(minpd|maxpd)$"), 2),
    ("cmp", re.compile(r"^(u?comisd|cmp\w*sd)$"), 1),
    ("cvt", re.compile(r"^cvt\w*(sd|si)\w*$"), 1),
    ("signbit", re.compile(r"^(andpd|andnpd|orpd|xorpd)$"), 1),
    ("f32", re.compile(r"^(add|sub|mul|div|sqrt)(ss|ps)$"), 1),
]
FLOP_CLASSES = ("add", "mul", "div", "sqrt", "minmax")          # what F_step sums
LIBM = re.compile(r"<(sin|cos|sincos|tan|asin|acos|atan|atan2|pow|exp|log|sqrt|cbrt|hypot|fmod|floor|ceil|round|trunc|modf|"
                  r"fabs|fmin|fmax|copysign)@plt>")

GROUPS = [
    ("driver (not part of the path)", re.compile(r"^(ref_|\(anonymous namespace\)|par_for|std::thread|void std::|std::)")),
    ("roots: companion matrix + eigen-solve (Eigen shim)", re.compile(r"Eigen::(EigenSolver|shim::(make_householder|house_left|house_right|hessenberg|real_schur_from_hessenberg|"
                                                                        r"eigenvalues_general))|long_term_planner::(roots|getSmallestPositiveNonComplexRoot)")),
    ("LongTermPlanner (optSwitchTimes, timeScaling, getTrajectory, ...)", re.compile(r"long_term_planner::")),
    ("capsule distance + Point algebra (ReachLib intersections / point)", re.compile(r"^(occupancy_containers::intersections::|point::)")),
    ("human reach models (ReachLib occupancies / systems / containers)", re.compile(r"^(occupancies::|occupancy_containers::|obstacles::|systems::)")),
    ("HumanReach (measurement, analysis)", re.compile(r"safety_shield::HumanReach")),
    ("RobotReach (FK, capsule transform, PFL velocities)", re.compile(r"safety_shield::RobotReach")),
    ("VerifyISO", re.compile(r"safety_shield::VerifyISO")),
    ("LongTermTraj (interpolate, windows, PFL ctor)", re.compile(r"safety_shield::LongTermTraj")),
    ("Path / Motion", re.compile(r"safety_shield::(Path|Motion)")),
    ("SafetyShield (step, planSafetyShield, calculateLongTermTrajectory, ...)", re.compile(r"safety_shield::SafetyShield")),
    ("Eigen shim: dense matrix / vector products", re.compile(r"Eigen::")),
]

# functions whose number of calls describes the workload
PROBES = {
    "SafetyShield::step": re.compile(r"^safety_shield::SafetyShield::step\("),
    "calculateLongTermTrajectory (replans)": re.compile(r"^safety_shield::SafetyShield::calculateLongTermTrajectory\("),
    "LongTermPlanner::planTrajectory": re.compile(r"^long_term_planner::LongTermPlanner::planTrajectory\("),
    "optSwitchTimes": re.compile(r"^long_term_planner::LongTermPlanner::optSwitchTimes\("),
    "timeScaling": re.compile(r"^long_term_planner::LongTermPlanner::timeScaling\("),
    "verify_human_reach": re.compile(r"^safety_shield::VerifyISO::verify_human_reach\("),
    "capsule_capsule distance (min_segment_distance)": re.compile(r"^occupancy_containers::intersections::min_segment_distance\("),
    "LongTermTraj::interpolate": re.compile(r"^safety_shield::LongTermTraj::interpolate\("),
}


def parse_disassembly(lib):
    """-> blocks: {return address: {"fn": mangled, "ops": Counter, "libm": Counter}}, entry: {mangled: first return address}."""
    txt = subprocess.run(["objdump", "-d", "--no-show-raw-insn", lib], check=True, capture_output=True, text=True).stdout
    fn_re = re.compile(r"^[0-9a-f]+ <(.+)>:$")
    ins_re = re.compile(r"^\s*([0-9a-f]+):\t(\S+)\s*(.*)$")
    blocks, entry = {}, {}
    fn, cur, pending_ret, head = None, None, False, None
    for line in txt.splitlines():
        m = fn_re.match(line)
        if m:
            fn, cur, pending_ret = m.group(1), None, False
            head = {"fn": fn, "ops": collections.Counter(), "libm": collections.Counter()}   # code ahead of the first call
            continue
        m = ins_re.match(line)
        if not m or fn is None:
            continue
        addr, mnem, rest = int(m.group(1), 16), m.group(2), m.group(3)
        if pending_ret:                      # the instruction after a trace call: its address is the block's key
            pending_ret = False
            cur = blocks.setdefault(addr, {"fn": fn, "ops": collections.Counter(), "libm": collections.Counter()})
            if head is not None:             # prologue of the function belongs to its first block
                cur["ops"].update(head["ops"])
                cur["libm"].update(head["libm"])
                entry.setdefault(fn, addr)
                head = None
        if mnem == "call" and "__sanitizer_cov_trace_pc" in rest:
            pending_ret = True
            continue
        tgt = cur if cur is not None else head
        if tgt is None:
            continue
        if mnem in ("call", "jmp"):
            lm = LIBM.search(rest)
            if lm:
                tgt["libm"][lm.group(1)] += 1
            continue
        for name, rx, w in CLASSES:
            if rx.match(mnem):
                tgt["ops"][name] += w
                break
    return blocks, entry


def bind(L):
    L.cov_dump.restype = C.c_long
    L.cov_dump.argtypes = [C.c_int, C.c_void_p, C.c_long]
    return L


def read_phase(L, phase):
    """(return address - load base, executions) rows of one phase."""
    cnt = L.cov_dump(phase, None, 0)
    buf = np.zeros((max(cnt, 1), 2), np.uint64)
    L.cov_dump(phase, buf.ctypes.data, cnt)
    return buf[:cnt]


def phase_totals(L, blocks, phase):
    """Operation class -> executed instructions of one phase (every function of the library)."""
    tot = collections.Counter()
    for pc, c in read_phase(L, phase):
        b = blocks.get(int(pc))
        if b is not None:
            for k, v in b["ops"].items():
                tot[k] += v * int(c)
            for k, v in b["libm"].items():
                tot["libm:" + k] += v * int(c)
    return tot


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.split("\n")
    return dict(zip(names, out))


def group_of(dem):
    for g, rx in GROUPS:
        if rx.search(dem):
            return g
    return "other"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--robot", default="schunk")
    ap.add_argument("--mocap", default="mujoco_mocap")
    ap.add_argument("--shield", default="ssm")
    ap.add_argument("--envs", type=int, default=2048)
    ap.add_argument("--warmup", type=int, default=25)
    ap.add_argument("--ticks", type=int, default=100)
    ap.add_argument("--traj", default=None)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    if not os.path.exists(COV_LIB):
        sys.exit("oracle/_ref/libsara_ref_cov.so is missing: make -C oracle flopcount (needs /root/reference)")

    import bench
    from oracle import ref
    ref.LIB_PATH = COV_LIB                                          # before the first ref.lib()
    L = bind(ref.lib())

    bargs = bench.parse_args(["--envs", str(a.envs), "--robot", a.robot, "--mocap", a.mocap, "--shield", a.shield]
                             + (["--traj", a.traj] if a.traj else []))
    files, sc, names = _files(bench, bargs)
    n = a.envs
    scen, kp, goals, masks = bench.make_inputs(bargs, sc, n, bargs.seed + 7919, names)
    chk = ref.RefBatch(n, bench.SAMPLE_TIME, files[0], files[1], files[2], np.zeros((n, 6)), scen.init_q,
                       shield_type=2 if a.shield == "pfl" else 1, threads=1)
    mb = masks.astype(bool)
    branch = np.zeros(3, np.int64)
    for k in range(a.warmup + a.ticks):
        t = k * bench.SAMPLE_TIME
        L.cov_enable(int(k >= a.warmup))
        L.cov_phase(0)
        chk.human_measurement(kp[bench.ring_index(k, bargs.ring)], t)
        L.cov_phase(1)
        chk.new_goal(goals[(k // bench.GOAL_PERIOD) % goals.shape[0]], mask=mb[k % bench.GOAL_PERIOD])
        L.cov_phase(2)
        o = chk.step(t)
        if k >= a.warmup:
            branch += np.bincount(o["branch"], minlength=3)[:3]
    L.cov_enable(0)

    blocks, entry = parse_disassembly(COV_LIB)
    dem = demangle(sorted({b["fn"] for b in blocks.values()}))
    steps = n * a.ticks
    per_phase, per_group, per_fn, libm, probes = {}, collections.defaultdict(collections.Counter), collections.Counter(), \
        collections.Counter(), collections.Counter()
    unknown = 0
    for ph, ph_name in enumerate(PHASES):
        tot = collections.Counter()
        for pc, c in read_phase(L, ph):
            b = blocks.get(int(pc))
            if b is None:
                unknown += 1
                continue
            c = int(c)
            d = dem[b["fn"]]
            g = group_of(d)
            for k2, v in b["ops"].items():
                tot[k2] += v * c
                per_group[g][k2] += v * c
                if k2 in FLOP_CLASSES:
                    per_fn[d] += v * c
            for k2, v in b["libm"].items():
                libm[k2] += v * c
                per_group[g]["libm_calls"] += v * c
            if entry.get(b["fn"]) == int(pc):
                for pn, rx in PROBES.items():
                    if rx.match(d):
                        probes[pn] += c
        per_phase[ph_name] = {k2: tot[k2] / steps for k2 in sorted(tot)}

    def flops(cn):
        return sum(cn.get(k2, 0) for k2 in FLOP_CLASSES)

    drv = "driver (not part of the path)"
    total = collections.Counter()
    for g, cn in per_group.items():
        if g != drv:
            total.update(cn)
    res = {
        "what": "FP64 operations per shield-step of the unmodified reference (oracle/_ref/libsara_ref_cov.so), block counts x "
                "objdump; F_step = add + sub + mul + div + sqrt + min/max instructions; compares, conversions, sign-bit "
                "operations and libm calls listed beside it",
        "workload": f"{a.robot} {a.shield.upper()} {a.mocap} human{(", " + a.traj) if a.traj else ""}, {n} envs x {a.ticks} ticks after {a.warmup} warm-up "
                    f"ticks of bench.make_inputs (goal period {bench.GOAL_PERIOD})",
        "shield_steps": steps,
        "branch_hist": branch.tolist(),
        "calls_per_step": {k2: v / steps for k2, v in sorted(probes.items())},
        "F_step": flops(total) / steps,
        "per_step": {k2: total[k2] / steps for k2 in sorted(total)},
        "libm_calls_per_step": {k2: v / steps for k2, v in sorted(libm.items())},
        "per_phase": {p: dict(v, flop=sum(v.get(k2, 0) for k2 in FLOP_CLASSES)) for p, v in per_phase.items()},
        "per_group": {g: dict({k2: cn[k2] / steps for k2 in sorted(cn)}, flop=flops(cn) / steps)
                      for g, cn in sorted(per_group.items(), key=lambda kv: -flops(kv[1]))},
        "top_functions_flop_per_step": [[d[:140], v / steps] for d, v in per_fn.most_common(25)],
        "blocks_instrumented": len(blocks),
        "return_addresses_not_in_disassembly": unknown,
    }
    js = json.dumps(res, indent=1)
    print(js)
    if a.out:
        with open(a.out, "w") as f:
            f.write(js + "\n")


def _files(bench, bargs):
    from human_robot_gym_b200 import workloads
    from human_robot_gym_b200.config import ShieldConfig
    files = workloads.default_files(bargs.robot, bargs.mocap, bargs.traj)
    sc = ShieldConfig.from_files(bench.SAMPLE_TIME, *files)
    return files, sc, workloads.mocap_joint_names(files[2])


if __name__ == "__main__":
    main()
