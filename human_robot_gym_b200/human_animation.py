"""Animation -> keypoints on the device: the recorded human of human-robot-gym without MuJoCo.

Mirrors what `HumanEnv` does for the shield's human input: `_setup_references` (human_env.py:1462-1482, animation step
length and joint names), `_reset_internal` (:1640-1657, placement and animation list), `_control_human` (:1710-1767,
frame selection, Pelvis mocap pose, hinge angles) and `human_measurement` (:459-463, the 23 `Human_<element>` site
positions after `sim.forward()`), for N envs at once. The kinematic tree comes from the reference's own
HRG/models/assets/human/human.xml (`load_human_tree`), the motion from its `<clip>.pkl` / `<clip>_info.json` files
(`load_clip`). The keypoint tensor it returns is what `BatchedSafetyShield.humanMeasurement` / `tick_device` take.
"""
from __future__ import annotations

import ctypes as C
import json
import pickle
import xml.etree.ElementTree as ET
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import cabi
from .batched import _ptr


def load_human_tree(xml_path: str) -> dict:
    """Parse human.xml: bodies below the Pelvis in document order -> {names, parent, anchor}.

    Checks what the device kernel assumes about the model (and the reference's XML satisfies): body pos 0 / quat identity,
    per body exactly three hinges with axes z, y, x in that order, all at the position of the body's site.
    """
    root = ET.parse(xml_path).getroot()
    pelvis = next(b for b in root.iter("body") if b.get("name") == "Pelvis")
    names: List[str] = []
    parent: List[int] = []
    anchor: List[List[float]] = []

    def vec(s, default):
        return [float(x) for x in s.split()] if s is not None else default

    def walk(body, parent_idx):
        for child in body.findall("body"):
            name = child.get("name")
            if vec(child.get("pos"), [0, 0, 0]) != [0, 0, 0] or vec(child.get("quat"), [1, 0, 0, 0]) != [1, 0, 0, 0]:
                raise ValueError(f"{name}: body pos/quat other than identity is not supported")
            site = child.find("site")
            joints = child.findall("joint")
            axes = [vec(j.get("axis"), None) for j in joints]
            if [j.get("name") for j in joints] != [f"{name}_{a}" for a in "zyx"] or axes != [[0, 0, 1], [0, 1, 0], [1, 0, 0]]:
                raise ValueError(f"{name}: expected hinges {name}_z, _y, _x about the z, y, x axes")
            c = vec(site.get("pos"), [0, 0, 0])
            if any(vec(j.get("pos"), [0, 0, 0]) != c for j in joints) or any(j.get("type") != "hinge" for j in joints):
                raise ValueError(f"{name}: hinges must sit at the body's site")
            names.append(name)
            parent.append(parent_idx)
            anchor.append(c)
        # document order is depth-first; the reference's keypoint order (HumanObject.joint_elements) is re-established below
        idx = {n: i for i, n in enumerate(names)}
        for child in body.findall("body"):
            walk(child, idx[child.get("name")])

    walk(pelvis, -1)
    return {"names": names, "parent": parent, "anchor": anchor}


def reorder_tree(tree: dict, joint_elements: Sequence[str]) -> dict:
    """The same tree with its bodies in `joint_elements` order (HRG/models/objects/human/human.py:58-82 = the keypoint
    order of `human_measurement` and of SS/config/mujoco_mocap.yaml)."""
    old = {n: i for i, n in enumerate(tree["names"])}
    new = {n: i for i, n in enumerate(joint_elements)}
    if sorted(old) != sorted(new):
        raise ValueError("joint_elements and the XML tree name different bodies")
    parent = []
    for n in joint_elements:
        p = tree["parent"][old[n]]
        parent.append(-1 if p < 0 else new[tree["names"][p]])
    if any(p >= i for i, p in enumerate(parent)):
        raise ValueError("joint_elements must list parents before their children")
    return {"names": list(joint_elements), "parent": parent, "anchor": [tree["anchor"][old[n]] for n in joint_elements]}


# HumanObject._setup_joint_names (HRG/models/objects/human/human.py:58-82)
JOINT_ELEMENTS = ["L_Hip", "R_Hip", "Torso", "L_Knee", "R_Knee", "Spine", "L_Ankle", "R_Ankle", "Chest", "L_Toe", "R_Toe",
                  "Neck", "L_Thorax", "R_Thorax", "Head", "L_Shoulder", "R_Shoulder", "L_Elbow", "R_Elbow", "L_Wrist",
                  "R_Wrist", "L_Hand", "R_Hand"]


def load_clip(pkl_path: str, info_path: Optional[str] = None) -> Tuple[Dict[str, np.ndarray], dict]:
    """(animation dict, info dict) as `HumanEnv` holds them in `human_animation_data` (human_env.py:1486-1530)."""
    with open(pkl_path, "rb") as f:
        anim = pickle.load(f)
    with open(info_path or pkl_path[:-4] + "_info.json") as f:
        info = json.load(f)
    return {k: np.asarray(v, dtype=np.float64) for k, v in anim.items()}, info


def load_tree_json(path: str) -> dict:
    """A tree written by `json.dump(reorder_tree(load_human_tree(...)))` (tests/golden/make_human_fixture.py)."""
    with open(path) as f:
        return json.load(f)


def load_clips_npz(path: str):
    """Clips stored channel-packed in one .npz (keys<i>, chan<i>, quat<i>, info<i>; tests/golden/make_human_fixture.py)
    -> list of (animation dict, info dict) as `load_clip` returns them."""
    z = np.load(path)
    clips, i = [], 0
    while f"keys{i}" in z:
        keys = [k for k in z[f"keys{i}"].tolist() if k != "Pelvis_quat"]
        anim = {k: np.ascontiguousarray(z[f"chan{i}"][:, j]) for j, k in enumerate(keys)}
        anim["Pelvis_quat"] = np.ascontiguousarray(z[f"quat{i}"])
        info = {"position_offset": z[f"info{i}"][:3].tolist(), "orientation_quat": z[f"info{i}"][3:].tolist()}
        clips.append((anim, info))
        i += 1
    return clips


def pack_clip(anim: Dict[str, np.ndarray], names: Sequence[str]) -> np.ndarray:
    """[frames, 7 + 3K]: Pelvis_pos xyz, Pelvis_quat xyzw, then the x, y, z angle of every body of `names`."""
    cols = [anim["Pelvis_pos_x"], anim["Pelvis_pos_y"], anim["Pelvis_pos_z"]] + list(np.asarray(anim["Pelvis_quat"]).T)
    for n in names:
        cols += [anim[n + "_x"], anim[n + "_y"], anim[n + "_z"]]
    return np.ascontiguousarray(np.stack(cols, axis=1), dtype=np.float64)


class BatchedHumanAnimation:
    """N replaying humans on one GPU.

    tree: {names, parent, anchor} in keypoint order; clips: list of (animation dict, info dict);
    human_animation_freq / model_timestep as in `HumanEnv.__init__` (defaults 120 fps, 0.002 s);
    n_animations_per_env = `_n_animations_to_sample_at_resets`.
    """

    def __init__(self, tree: dict, clips, n_envs: int, device="cuda:0", human_animation_freq: float = 120.0,
                 model_timestep: float = 0.002, n_animations_per_env: int = 1):
        self.lib = cabi.load_library()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise cabi.ShieldLibraryError("BatchedHumanAnimation needs a CUDA device: there is no CPU fallback")
        self.tree, self.n_envs, self.K = tree, int(n_envs), len(tree["names"])
        self.list_len = int(n_animations_per_env)
        self.step_length = int(1 / model_timestep) / human_animation_freq  # human_env.py:1462-1465
        frames = [pack_clip(a, tree["names"]) for a, _ in clips]
        self.clip_len = np.array([f.shape[0] for f in frames], dtype=np.int32)
        bank = np.ascontiguousarray(np.concatenate(frames, axis=0))
        info = np.ascontiguousarray(
            [list(i["position_offset"]) + list(i["orientation_quat"]) for _, i in clips], dtype=np.float64)
        parent = np.ascontiguousarray(tree["parent"], dtype=np.int32)
        anchor = np.ascontiguousarray(tree["anchor"], dtype=np.float64)
        h = C.c_void_p()
        cabi.check(self.lib.hrg_human_create(self.K, parent.ctypes.data, anchor.ctypes.data, len(clips),
                                             self.clip_len.ctypes.data, bank.ctypes.data, info.ctypes.data,
                                             float(self.step_length), self.n_envs, self.list_len, self.device.index or 0,
                                             C.byref(h)), "hrg_human_create")
        self._h = h
        self.keypoints = torch.zeros(self.n_envs, self.K, 3, dtype=torch.float64, device=self.device)
        self.clip = torch.zeros(self.n_envs, dtype=torch.int32, device=self.device)
        self.frame = torch.zeros(self.n_envs, dtype=torch.int32, device=self.device)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def reset(self, human_pos_offset, human_rot_offset, animation_ids, env_mask=None):
        """`_reset_internal`: human_pos_offset [N,3], human_rot_offset [N,4] (w,x,y,z), animation_ids [N,list_len]."""
        f64 = dict(dtype=torch.float64, device=self.device)
        place = torch.cat([torch.as_tensor(human_pos_offset, **f64).reshape(self.n_envs, 3),
                           torch.as_tensor(human_rot_offset, **f64).reshape(self.n_envs, 4)], dim=1).contiguous()
        ids = torch.as_tensor(animation_ids, dtype=torch.int32, device=self.device).reshape(self.n_envs, self.list_len)
        if int(ids.min()) < 0 or int(ids.max()) >= len(self.clip_len):
            raise ValueError("animation id out of range")
        m = None if env_mask is None else torch.as_tensor(env_mask, dtype=torch.uint8, device=self.device).contiguous()
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_human_reset(self._h, _ptr(m), _ptr(place), _ptr(ids.contiguous()), self._stream()),
                       "hrg_human_reset")

    def human_measurement(self, low_level_time, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`_control_human(force_update=True)` + `human_measurement` at low_level_time (int, or [N] int64 tensor)."""
        if not torch.is_tensor(low_level_time):
            low_level_time = torch.full((self.n_envs,), int(low_level_time), dtype=torch.int64, device=self.device)
        t = low_level_time.to(device=self.device, dtype=torch.int64).contiguous()
        kp = self.keypoints if out is None else out
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_human_keypoints(self._h, _ptr(t), _ptr(kp), _ptr(self.clip), _ptr(self.frame),
                                                    self._stream()), "hrg_human_keypoints")
        return kp

    def tick_replay_host(self, shield, low_level_time_host, t_host, goal_mask_host, goal_q_host, q_host, dq_host, ddq_host,
                         flags_host):
        """One control tick with page-locked HOST tensors when the human is this replay: clock + goals up, keypoints made and
        consumed on the device, setpoints down (hrg_shield_tick_replay_host). The caller synchronises the stream."""
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_shield_tick_replay_host(
                shield._h, self._h, _ptr(low_level_time_host), _ptr(t_host), _ptr(goal_mask_host), _ptr(goal_q_host),
                _ptr(q_host), _ptr(dq_host), _ptr(ddq_host), _ptr(flags_host), self._stream()), "hrg_shield_tick_replay_host")

    def tick_replay_piped(self, shield, slot, low_level_time_host, t_host, goal_mask_host, goal_q_host, q_host, dq_host,
                          ddq_host, flags_host, copy_stream):
        """tick_replay_host with the setpoints leaving on `copy_stream` (a torch.cuda.Stream) out of output slot 0/1, so that
        the device->host copy of this tick overlaps the kernels of the next (hrg_shield_tick_replay_piped). Alternate the
        slots, give each slot its own host buffers (inputs too), and wait for an event recorded on `copy_stream` after the
        call before reading the slot's setpoints."""
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_shield_tick_replay_piped(
                shield._h, self._h, int(slot), _ptr(low_level_time_host), _ptr(t_host), _ptr(goal_mask_host), _ptr(goal_q_host),
                _ptr(q_host), _ptr(dq_host), _ptr(ddq_host), _ptr(flags_host), self._stream(), copy_stream.cuda_stream),
                "hrg_shield_tick_replay_piped")

    @property
    def launch_count(self) -> int:
        return int(self.lib.hrg_human_launch_count(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self.lib.hrg_human_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
