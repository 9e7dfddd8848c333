"""The reference's OWN binding tests, collected UNCHANGED (SS/tests/bindings/test_motion.py, test_ltt.py,
test_safety_shield.py), against the top-level `safety_shield_py` package of this repository.

They `from safety_shield_py import ...` and find their YAML files through "<their dir>/../../config". Where /root/reference
is mounted they are collected from there; on the GPU box from the unmodified copies oracle/Makefile stages under
oracle/_ref/safety_shield (git-ignored, never part of this repository's history). Each file runs in its own pytest process with this
repository root on PYTHONPATH, exactly like a user of the reference would run them after switching modules."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SS = "human_robot_gym/controllers/failsafe_controller/sara-shield/safety_shield"
_CANDIDATES = [os.path.join("/root/reference", _SS, "tests", "bindings"),
               os.path.join(ROOT, "oracle", "_ref", "safety_shield", "tests", "bindings")]


def _bindings_dir():
    for d in _CANDIDATES:
        if os.path.isfile(os.path.join(d, "test_safety_shield.py")):
            return d
    pytest.skip("reference binding tests not available (neither /root/reference nor oracle/_ref/safety_shield)")


def _run(name):
    d = _bindings_dir()
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    # /root/reference is read-only: no cache dir, no bytecode next to the collected file
    env["PYTHONDONTWRITEBYTECODE"] = "1"
    r = subprocess.run([sys.executable, "-m", "pytest", "-p", "no:cacheprovider", "-q", "-x", "--rootdir", d,
                        os.path.join(d, name)], env=env, cwd=ROOT, capture_output=True, text=True, timeout=900)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "failed" not in r.stdout, tail
    return r.stdout


def test_reference_test_motion_unchanged():
    out = _run("test_motion.py")
    print(out[-200:])


def test_reference_test_ltt_unchanged():
    out = _run("test_ltt.py")
    print(out[-200:])


@pytest.mark.gpu
def test_reference_test_safety_shield_unchanged():
    """Constructs shields from the reference's modrob1 / Schunk YAML files, steps them on the GPU and compares the golden
    robot capsules of test_safety_shield.py:151-290."""
    out = _run("test_safety_shield.py")
    print(out[-200:])
