"""Host-side parsing of the reference's three YAML config files into flat constant tables.

Mirrors what the reference constructor reads (SS/src/safety_shield.cc:61-131): robot kinematics + capsules
(`robot_parameters_*.yaml`), limits (`trajectory_parameters_*.yaml`) and the human model (`*_mocap.yaml`).
The reference parses with yaml-cpp, which is more lenient than PyYAML in two ways that the shipped files rely on:
tab characters in flow sequences (`robot_parameters_ur5e.yaml:9`) and exponent floats without a dot (`1e-15`).
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np
import yaml

_FLOAT_RE = re.compile(r"^[-+]?(\d+\.?\d*|\.\d+)([eE][-+]?\d+)?$")


def _fix_numbers(x):
    if isinstance(x, dict):
        return {k: _fix_numbers(v) for k, v in x.items()}
    if isinstance(x, list):
        return [_fix_numbers(v) for v in x]
    if isinstance(x, str) and _FLOAT_RE.match(x.strip()):
        return float(x)
    return x


def load_yaml(path: str) -> dict:
    """`YAML::LoadFile` equivalent (SS/src/safety_shield.cc:61,70,97); raises like the reference on a missing file."""
    with open(path) as f:
        text = f.read().replace("\t", " ")
    return _fix_numbers(yaml.safe_load(text))


@dataclass
class RobotConfig:
    """`robot_parameters_*.yaml` → RobotReach ctor arguments (SS/src/robot_reach.cc:5-39)."""
    name: str
    nb_joints: int
    transformation_matrices: np.ndarray  # [J, 16] row-major 4x4 fixed transforms
    enclosures: np.ndarray               # [J, 7]  p1(3) p2(3) r   in link frame
    secure_radius: float

    @staticmethod
    def from_dict(d: dict) -> "RobotConfig":
        J = int(d["nb_joints"])
        tm = np.asarray(d["transformation_matrices"], dtype=np.float64)
        en = np.asarray(d["enclosures"], dtype=np.float64)
        if tm.size < 16 * J or en.size < 7 * J:
            raise ValueError("robot config: transformation_matrices/enclosures shorter than nb_joints")
        return RobotConfig(str(d["robot_name"]), J, tm[: 16 * J].reshape(J, 16).copy(), en[: 7 * J].reshape(J, 7).copy(),
                           float(d["secure_radius"]))


@dataclass
class TrajectoryConfig:
    """`trajectory_parameters_*.yaml` (SS/src/safety_shield.cc:70-86)."""
    max_s_stop: float
    q_min_allowed: np.ndarray
    q_max_allowed: np.ndarray
    v_max_allowed: np.ndarray
    a_max_allowed: np.ndarray
    j_max_allowed: np.ndarray
    a_max_ltt: np.ndarray
    j_max_ltt: np.ndarray
    v_safe: float
    alpha_i_max: float
    velocity_method: int

    @staticmethod
    def from_dict(d: dict) -> "TrajectoryConfig":
        arr = lambda k: np.asarray(d[k], dtype=np.float64)
        return TrajectoryConfig(float(d["max_s_stop"]), arr("q_min_allowed"), arr("q_max_allowed"), arr("v_max_allowed"),
                                arr("a_max_allowed"), arr("j_max_allowed"), arr("a_max_ltt"), arr("j_max_ltt"),
                                float(d["v_safe"]), float(d["alpha_i_max"]), int(d["velocity_method"]))


@dataclass
class HumanConfig:
    """Human model tables in the order the reference EVALUATES them.

    The reference stores bodies/extremities in `std::map<std::string, ...>`, so capsules come out in lexicographic
    order of the body name (RL/src/articulated_vel.cpp:38-41), not YAML order. `ArticulatedPos` additionally indexes
    its thickness/length/max_v vectors (filled in YAML order, SS/src/human_reach.cc:28-36) by that sorted iteration
    index (RL/src/articulated_pos.cpp:42-46); the same mismatch is reproduced here.
    """
    n_keypoints: int
    measurement_error_pos: float
    measurement_error_vel: float
    delay: float
    joint_v_max: np.ndarray
    joint_a_max: np.ndarray
    body_names: List[str]
    body_prox: np.ndarray      # [B] int32
    body_dist: np.ndarray      # [B] int32
    body_thickness: np.ndarray  # [B]
    ext_names: List[str]
    ext_base: np.ndarray       # [E] int32
    ext_thickness: np.ndarray  # [E]
    ext_length: np.ndarray     # [E]
    ext_v_max: np.ndarray      # [E]

    @property
    def n_bodies(self) -> int:
        return len(self.body_names)

    @property
    def n_ext(self) -> int:
        return len(self.ext_names)

    @property
    def n_capsules(self) -> int:
        return self.n_ext + 2 * self.n_bodies

    @staticmethod
    def from_dict(d: dict) -> "HumanConfig":
        names = list(d["joint_names"])
        idx: Dict[str, int] = {}
        for i, n in enumerate(names):
            idx[n] = i  # later duplicates overwrite, like std::map::operator[] (safety_shield.cc:104-106)
        vmax = np.asarray(d["joint_v_max"], dtype=np.float64)
        amax = np.asarray(d["joint_a_max"], dtype=np.float64)
        bodies: Dict[str, Tuple[int, int, float]] = {}
        for b in d.get("bodies") or []:
            bodies[str(b["name"])] = (idx[b["proximal"]], idx[b["distal"]], float(b["thickness"]))
        bn = sorted(bodies)  # std::map iteration order == byte-wise lexicographic
        ext_yaml = d.get("extremities") or []
        e_base = [str(e["base"]) for e in ext_yaml]
        e_end = [str(e["end"]) for e in ext_yaml]
        e_len = [float(e["length"]) for e in ext_yaml]
        e_thk = [float(e["thickness"]) for e in ext_yaml]
        e_vmx = [max(vmax[idx[b]], vmax[idx[e]]) for b, e in zip(e_base, e_end)]
        emap = {b: idx[b] for b in e_base}  # keyed by BASE name (human_reach.cc:30)
        en = sorted(emap)
        ne = len(en)
        return HumanConfig(
            n_keypoints=len(idx), measurement_error_pos=float(d["measurement_error_pos"]),
            measurement_error_vel=float(d["measurement_error_vel"]), delay=float(d["delay"]),
            joint_v_max=vmax, joint_a_max=amax, body_names=bn,
            body_prox=np.asarray([bodies[n][0] for n in bn], dtype=np.int32),
            body_dist=np.asarray([bodies[n][1] for n in bn], dtype=np.int32),
            body_thickness=np.asarray([bodies[n][2] for n in bn], dtype=np.float64),
            ext_names=en, ext_base=np.asarray([emap[n] for n in en], dtype=np.int32),
            # sorted iteration index i picks the i-th YAML-order entry
            ext_thickness=np.asarray(e_thk[:ne], dtype=np.float64), ext_length=np.asarray(e_len[:ne], dtype=np.float64),
            ext_v_max=np.asarray(e_vmx[:ne], dtype=np.float64))


@dataclass
class ShieldConfig:
    sample_time: float
    robot: RobotConfig
    traj: TrajectoryConfig
    human: HumanConfig
    files: Tuple[str, str, str] = field(default=("", "", ""))

    @staticmethod
    def from_files(sample_time: float, trajectory_config_file: str, robot_config_file: str,
                   mocap_config_file: str) -> "ShieldConfig":
        return ShieldConfig(float(sample_time), RobotConfig.from_dict(load_yaml(robot_config_file)),
                            TrajectoryConfig.from_dict(load_yaml(trajectory_config_file)),
                            HumanConfig.from_dict(load_yaml(mocap_config_file)),
                            (trajectory_config_file, robot_config_file, mocap_config_file))
