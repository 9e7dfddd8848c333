"""CPU: the C-ABI library loads, exports every symbol include/hrg_shield.h declares, its structs match the ctypes
mirrors, and the product path fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from human_robot_gym_b200 import cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "hrg_shield.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hrg_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    names = declared_functions()
    # one entry point per method of the reference's pybind class (SS/bindings/safety_shield_bindings.cpp:82-115)
    for must in ("hrg_shield_create", "hrg_shield_reset", "hrg_shield_step", "hrg_shield_new_goal", "hrg_shield_set_ltt",
                 "hrg_shield_human_measurement", "hrg_shield_get_capsules", "hrg_shield_destroy"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = cabi.load_library()
    names = declared_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hrg_shield.h but not exported by libhrg_shield.so"
        assert n in cabi.SYMBOLS, f"{n} has no ctypes prototype in cabi.SYMBOLS"
    assert sorted(cabi.SYMBOLS) == names


def test_struct_layout_matches_the_header(tmp_path):
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "hrg_shield.h"\n'
                    'int main(void){printf("%zu %zu %zu %zu %zu\\n", sizeof(HrgConfig), sizeof(HrgStats), '
                    'offsetof(HrgConfig, link_T), offsetof(HrgConfig, ext_vmax), offsetof(HrgStats, min_dist));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(prog)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert out[0] == C.sizeof(cabi.HrgConfig)
    assert out[1] == C.sizeof(cabi.HrgStats)
    assert out[2] == cabi.HrgConfig.link_T.offset
    assert out[3] == cabi.HrgConfig.ext_vmax.offset
    assert out[4] == cabi.HrgStats.min_dist.offset


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_cuda(), reason="checks the no-device error path")
def test_no_cpu_fallback(config_dir):
    """Without a CUDA device creation fails with HRG_ENODEVICE and the host class raises; nothing computes on the CPU."""
    from human_robot_gym_b200.config import ShieldConfig
    lib = cabi.load_library()
    d = config_dir
    sc = ShieldConfig.from_files(0.004, f"{d}/trajectory_parameters_schunk.yaml", f"{d}/robot_parameters_schunk.yaml",
                                 f"{d}/mujoco_mocap.yaml")
    cfg = cabi.make_cfg(sc, 64)
    pose, q0 = np.zeros((4, 6)), np.zeros((4, 6))
    h = C.c_void_p()
    rc = lib.hrg_shield_create(C.byref(cfg), 4, 0, pose.ctypes.data, q0.ctypes.data, None, C.byref(h))
    assert rc == cabi.ENODEVICE
    assert not h.value
    assert b"no CUDA device" in lib.hrg_last_error()
    from human_robot_gym_b200.batched import BatchedSafetyShield
    with pytest.raises(cabi.ShieldLibraryError):
        BatchedSafetyShield(4, 0.004, f"{d}/trajectory_parameters_schunk.yaml", f"{d}/robot_parameters_schunk.yaml",
                            f"{d}/mujoco_mocap.yaml")


def test_argument_validation_without_device():
    lib = cabi.load_library()
    assert lib.hrg_shield_create(None, 1, 0, None, None, None, C.byref(C.c_void_p())) == cabi.EINVAL
    assert lib.hrg_shield_step(None, None, None, None, None, None, None, None, None, None) == cabi.EINVAL
    assert lib.hrg_shield_num_envs(None) == 0
    assert lib.hrg_stage_capsule_dist(-1, None, None, None, None) == cabi.EINVAL


@pytest.mark.skipif(_has_cuda(), reason="checks the no-device error path")
def test_human_replay_has_no_cpu_fallback_either():
    """hrg_human_create validates its arguments, then fails with HRG_ENODEVICE without a GPU: nothing is computed on the host."""
    lib = cabi.load_library()
    parent = np.array([-1, 0], dtype=np.int32)
    anchor, frames = np.zeros((2, 3)), np.zeros((3, 13))
    frames[:, 6] = 1.0
    clip_len, info = np.array([3], dtype=np.int32), np.array([[0, 0, 0, 0, 0, 0, 1.0]])
    h = C.c_void_p()

    def create(n_bodies, step):
        return lib.hrg_human_create(n_bodies, parent.ctypes.data, anchor.ctypes.data, 1, clip_len.ctypes.data,
                                    frames.ctypes.data, info.ctypes.data, step, 4, 1, 0, C.byref(h))
    assert create(2, 500 / 120.0) == cabi.ENODEVICE and not h.value
    assert b"no CPU fallback" in lib.hrg_last_error()
    assert create(0, 500 / 120.0) == cabi.EINVAL       # n_bodies out of range
    assert create(2, 0.5) == cabi.EINVAL               # animation faster than the simulation (human_env.py:1466)
    assert lib.hrg_human_keypoints(None, None, None, None, None, None) == cabi.EINVAL
    assert lib.hrg_capsule_markers(-1, None, None, None, None, None) == cabi.EINVAL


def test_product_never_touches_the_oracle():
    """oracle/ and tests/ are test infrastructure: the package and the importable drop-in never import them, no CUDA / C++
    source of the library includes anything from there, bench.py executes oracle/ in its CPU leg only (cpu_loop: the
    `cpu_baseline` key and --impl reference) and __graft_entry__ in build() (compiling the checker) and smoke() (checking)."""
    import ast
    import glob
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def oracle_imports(path):
        """{enclosing top-level function or '<module>': [imported module names that start with oracle / tests]}"""
        with open(path) as f:
            tree = ast.parse(f.read())
        found = {}

        def visit(node, where):
            for ch in ast.iter_child_nodes(node):
                w = ch.name if where == "<module>" and isinstance(ch, (ast.FunctionDef, ast.ClassDef)) else where
                names = []
                if isinstance(ch, ast.Import):
                    names = [a.name for a in ch.names]
                elif isinstance(ch, ast.ImportFrom) and ch.module and ch.level == 0:
                    names = [ch.module]
                for nm in names:
                    if nm.split(".")[0] in ("oracle", "tests"):
                        found.setdefault(w, []).append(nm)
                visit(ch, w)
        visit(tree, "<module>")
        return found

    product = glob.glob(os.path.join(root, "human_robot_gym_b200", "**", "*.py"), recursive=True) \
        + glob.glob(os.path.join(root, "safety_shield_py", "**", "*.py"), recursive=True)
    assert len(product) > 5
    for p in product:
        assert oracle_imports(p) == {}, p
    for p in glob.glob(os.path.join(root, "human_robot_gym_b200", "csrc", "*")) + glob.glob(os.path.join(root, "include", "*")):
        with open(p, errors="replace") as f:
            for line in f:
                if line.lstrip().startswith("#include"):
                    assert "oracle" not in line and "tests/" not in line and "host_emul" not in line, (p, line)
    b = oracle_imports(os.path.join(root, "bench.py"))
    assert set(b) <= {"cpu_loop"} and all(n.startswith("oracle") for n in b.get("cpu_loop", [])), b
    g = oracle_imports(os.path.join(root, "__graft_entry__.py"))
    assert set(g) <= {"smoke", "build"}, g


def test_integration_doc_names_every_entry_point():
    """INTEGRATION.md maps each function include/hrg_shield.h declares to the reference interface it replaces (or says it
    has none): a new entry point must not ship without its row."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, "include", "hrg_shield.h")) as f:
        names = sorted(set(re.findall(r"\b(hrg_[a-z0-9_]+)\s*\(", f.read())))
    with open(os.path.join(root, "INTEGRATION.md")) as f:
        doc = f.read()
    wild = [m[:-1] for m in re.findall(r"`(hrg_[a-z0-9_]+\*)`", doc)]          # rows such as `hrg_stage_*`
    missing = [n for n in names if n not in doc and not any(n.startswith(w) for w in wild)]
    assert len(names) >= 40 and not missing, missing
