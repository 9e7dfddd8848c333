"""The benchmark / parity workloads' configuration files.

`data/shield_configs.json` holds the PARSED content of the reference's SS/config/*.yaml (numbers only; written by
tests/golden/make_configs.py in the build container, /root/reference does not exist on the GPU box). `config_dir()`
writes them back out as YAML files in the reference schema, so every consumer goes through the same constructor
arguments as a user of the reference (`SafetyShield(sample_time, trajectory_config_file, robot_config_file,
mocap_config_file, ...)`), plus the two builder-supplied variants BASELINE config 4 names ("full-body human capsule
set, long verification horizon").
"""
from __future__ import annotations

import json
import os
import tempfile

import yaml

_HERE = os.path.dirname(os.path.abspath(__file__))
CONFIGS_JSON = os.path.join(_HERE, "data", "shield_configs.json")
_CFG_DIR = None


def configs() -> dict:
    with open(CONFIGS_JSON) as f:
        return json.load(f)


def full_body_mocap(base: dict) -> dict:
    """mujoco_mocap.yaml + legs, pelvis and spine: 16 bodies, 4 extremities -> H = 4 + 16 + 16 = 36 capsules."""
    m = json.loads(json.dumps(base))
    extra = [("L_UpperLeg", "L_Hip", "L_Knee", 0.2), ("R_UpperLeg", "R_Hip", "R_Knee", 0.2),
             ("L_LowerLeg", "L_Knee", "L_Ankle", 0.132), ("R_LowerLeg", "R_Knee", "R_Ankle", 0.132),
             ("L_Foot", "L_Ankle", "L_Toe", 0.1), ("R_Foot", "R_Ankle", "R_Toe", 0.1),
             ("Pelvis", "L_Hip", "R_Hip", 0.3), ("Spine", "Torso", "Chest", 0.434)]
    for name, p, d, th in extra:
        m["bodies"].append({"name": name, "proximal": p, "distal": d, "thickness": th})
    m["extremities"].append({"name": "L_Leg", "base": "L_Hip", "end": "L_Toe", "length": 1.0, "thickness": 0.0})
    m["extremities"].append({"name": "R_Leg", "base": "R_Hip", "end": "R_Toe", "length": 1.0, "thickness": 0.0})
    return m


def config_dir() -> str:
    """The configs written back out as YAML files in the reference schema (once per process)."""
    global _CFG_DIR
    if _CFG_DIR is None:
        d = tempfile.mkdtemp(prefix="shield_cfg_")
        cfgs = configs()
        for name, content in cfgs.items():
            with open(os.path.join(d, name + ".yaml"), "w") as f:
                yaml.safe_dump(content, f, default_flow_style=None)
        with open(os.path.join(d, "mujoco_mocap_full_body.yaml"), "w") as f:
            yaml.safe_dump(full_body_mocap(cfgs["mujoco_mocap"]), f, default_flow_style=None)
        long_h = dict(cfgs["trajectory_parameters_schunk"])
        long_h["max_s_stop"] = 0.5
        with open(os.path.join(d, "trajectory_parameters_schunk_long.yaml"), "w") as f:
            yaml.safe_dump(long_h, f, default_flow_style=None)
        _CFG_DIR = d
    return _CFG_DIR


def default_files(robot="schunk", mocap="mujoco_mocap", traj=None):
    """(trajectory yaml, robot yaml, mocap yaml) paths — the reference constructor's three file arguments."""
    d = config_dir()
    traj = traj or f"trajectory_parameters_{robot}"
    return (os.path.join(d, traj + ".yaml"), os.path.join(d, f"robot_parameters_{robot}.yaml"),
            os.path.join(d, mocap + ".yaml"))


def mocap_joint_names(mocap_file: str):
    from .config import load_yaml
    return load_yaml(mocap_file)["joint_names"]
