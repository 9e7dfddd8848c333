"""Trust gate of the oracle: the reference's OWN gtest suites, compiled UNMODIFIED by
`make -C oracle gtest` (vendored googletest 1.11, SS/external/yaml-cpp/test/gtest-1.11.0) against the same objects and the
same oracle/eigen_shim as oracle/_ref/libsara_ref.so, must pass.

Eigen 3.4 is the one dependency of the reference that is not under /root/reference and not installed here; everything
the parity tests call "reference" therefore rests on the shim's fixed-size algebra, its dynamic matrices (the PFL
Jacobian / capsule-velocity code of SS/src/robot_reach.cc:110-199) and its real-Schur EigenSolver. These suites hold the
reference's known answers for exactly that code:

  robot_reach_tests    SS/tests/src/robot_reach_tests.cc:16-30,64-152 (FK, transformCapsule, reach), :182-508 (Jacobians,
                       capsule velocities: the whole PFL oracle)
  long_term_traj_tests :12-64 (index / clamp), :66-194 (sliding window), :196-230, :246-372 (interpolate and its edge cases),
                       :232-244 (PFL: approximate >= exact Cartesian velocity on Schunk)
  path_tests :21-66, motion_tests, verify_iso_tests :17-47, human_reach_tests :14-205, safety_shield_tests :16-53
                       (CalculateMaxAccJerkTest closed form, the 169 680-case planSafetyShield grid)
  long_term_planner_tests LTP/tests/src/long_term_planner_tests.cc:12-407 (optBraking, optSwitchTimes, timeScaling,
                       planTrajectory end positions, the (q, v, a) grids - every root solve goes through the shim's solver)
  roots_tests          LTP/tests/src/roots_tests.cc:9-32 (degree-6 roots<float>, which also pins the eigenvalue ORDER)

The binaries are built in the container that has /root/reference and travel to the GPU box prebuilt (oracle/_ref is
git-ignored, not gpurun-ignored); their fixtures load YAML relative to the working directory (`../config/...`,
`../../safety_shield/config/...`), so they run from oracle/_ref/safety_shield/build next to the staged config files."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "gtest")
CWD = os.path.join(ROOT, "oracle", "_ref", "safety_shield", "build")

# suite -> number of tests the reference's source holds (so that a suite that silently ran nothing fails)
SUITES = {"robot_reach_tests": 23, "long_term_traj_tests": 12, "path_tests": 3, "motion_tests": 6, "verify_iso_tests": 4,
          "human_reach_tests": 9, "safety_shield_tests": 3, "long_term_planner_tests": 10, "roots_tests": 1}
EXPECTED_FAILURES = {}   # none: every suite passes on the shim build, including roots<float>'s ordering


@pytest.mark.parametrize("suite", list(SUITES))
def test_reference_gtest_suite_passes_on_the_shim_build(suite):
    exe = os.path.join(BIN, suite)
    if not os.path.exists(exe) or not os.path.isdir(CWD):
        pytest.skip("oracle/_ref/gtest not built (run `make -C oracle gtest` where /root/reference exists)")
    r = subprocess.run([exe], cwd=CWD, capture_output=True, text=True, timeout=900)
    out = r.stdout + r.stderr
    m = re.search(r"\[  PASSED  \] (\d+) tests?\.", out)
    assert r.returncode == 0 and m, out[-3000:]
    assert "FAILED" not in out, out[-3000:]
    assert int(m.group(1)) == SUITES[suite], (suite, m.group(1))
