"""Regenerates human_robot_gym_b200/data/shield_configs.json from the reference's YAML config files (run in the build container only).

The reference ships robot kinematics/capsules, trajectory limits and human models as YAML under
SS/config/*.yaml. /root/reference does not exist on the GPU box, so the PARSED content
(numbers only, no comments/layout) is committed here as one JSON fixture; tests write it back out as YAML files in
a temp dir (tests/conftest.py) to exercise the reference-compatible constructor `SafetyShield(sample_time,
trajectory_config_file, robot_config_file, mocap_config_file, ...)`.
"""
import glob
import json
import os
import sys

import yaml  # noqa: F401

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from human_robot_gym_b200.config import load_yaml  # noqa: E402

REF = os.environ.get("HRG_REFERENCE", "/root/reference")
CFG = os.path.join(REF, "human_robot_gym/controllers/failsafe_controller/sara-shield/safety_shield/config")


def main():
    out = {}
    for path in sorted(glob.glob(os.path.join(CFG, "*.yaml"))):
        with open(path) as f:
            pass
        out[os.path.basename(path)[:-5]] = load_yaml(path)
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "human_robot_gym_b200", "data",
                       "shield_configs.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print("wrote", dst, len(out), "configs")


if __name__ == "__main__":
    sys.exit(main())
