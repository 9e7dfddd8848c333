"""TEST INFRASTRUCTURE ONLY - numpy restatement of the animation -> keypoints path.

PARITY UNPINNED for the kinematic tree: the reference obtains the keypoints from MuJoCo (`sim.forward()` +
`sim.data.get_site_xpos`, HRG/environments/manipulation/human_env.py:459-463), a third-party dependency that is absent
from /root/reference and from this image (mujoco-py 2.1 / MuJoCo 2.1 behind robosuite==1.3.2, reference
requirements.txt:8). `site_positions` restates MuJoCo's published `mj_kinematics` (engine_forward / engine_core_smooth.c:
body frame = parent frame * (pos, quat); per joint in XML order: anchor and axis into the world, hinge rotation composed
onto the body quaternion, body position corrected so that the anchor stays put; site = body frame * site pos) for the
tree of HRG/models/assets/human/human.xml. The root pose and the time management ARE the reference's own arithmetic
(`_control_human`, human_env.py:1710-1767, through scipy's Rotation exactly as the reference calls it).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import math

import numpy as np
from scipy.spatial.transform import Rotation


def _qmul(a, b):  # (w,x,y,z), mju_mulQuat
    return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
                     a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                     a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
                     a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def _qmat(q):  # mju_quat2Mat
    w, x, y, z = q
    return np.array([[w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def root_pose(anim, info, frame, human_pos_offset, human_rot_offset_wxyz):
    """Mocap position and (w,x,y,z) quaternion of the Pelvis - human_env.py:1737-1763, statement by statement."""
    human_base_quat = Rotation.from_quat([0.5, 0.5, 0.5, 0.5])  # human_env.py:373
    animation_pos = [anim["Pelvis_pos_x"][frame], anim["Pelvis_pos_y"][frame], anim["Pelvis_pos_z"][frame]]
    off = info["position_offset"]
    human_pos = [animation_pos[0] + off[0], animation_pos[1] + off[1], animation_pos[2] + off[2]]
    human_rot = human_base_quat * Rotation.from_quat(info["orientation_quat"])
    human_pos = human_rot.apply(human_pos)
    human_pos = [human_pos[i] + human_pos_offset[i] for i in range(3)]
    q = human_rot_offset_wxyz
    human_rot = Rotation.from_quat([q[1], q[2], q[3], q[0]]) * human_rot  # mjcf_utils.quat_to_rot
    human_rot = human_rot * Rotation.from_quat(anim["Pelvis_quat"][frame])
    xyzw = human_rot.as_quat()
    return np.asarray(human_pos, float), np.array([xyzw[3], xyzw[0], xyzw[1], xyzw[2]])  # rot_to_quat


def site_positions(tree, mocap_pos, mocap_quat, angles):
    """World positions [K,3] of the sites `Human_<joint_element>` for one frame.

    tree: dict(names, parent [K] (-1 = Pelvis), anchor [K,3]) - every body of human.xml has pos 0 / quat identity and three
    hinges z, y, x (XML order) at the site position. angles: dict name_{x,y,z} -> radians (qpos, qpos0 = 0).
    """
    K = len(tree["names"])
    quat0 = np.asarray(mocap_quat, float)
    quat0 = quat0 / np.linalg.norm(quat0)  # mj_kinematics normalises mocap_quat
    xpos = [None] * K
    xquat = [None] * K
    site = np.zeros((K, 3))
    axes = {"z": np.array([0.0, 0.0, 1.0]), "y": np.array([0.0, 1.0, 0.0]), "x": np.array([1.0, 0.0, 0.0])}
    for b in range(K):
        p = tree["parent"][b]
        ppos, pquat = (np.asarray(mocap_pos, float), quat0) if p < 0 else (xpos[p], xquat[p])
        pos, quat = ppos.copy(), pquat.copy()  # body pos 0, body quat identity
        jpos = np.asarray(tree["anchor"][b], float)
        for ax in ("z", "y", "x"):
            R = _qmat(quat)
            xanchor = R @ jpos + pos
            ang = angles[tree["names"][b] + "_" + ax]
            qloc = np.concatenate([[math.cos(ang / 2)], axes[ax] * math.sin(ang / 2)])  # mju_axisAngle2Quat
            quat = _qmul(quat, qloc)
            pos = xanchor - _qmat(quat) @ jpos  # keep the anchor where it was
        quat = quat / np.linalg.norm(quat)
        xpos[b], xquat[b] = pos, quat
        site[b] = pos + _qmat(quat) @ jpos  # site pos == joint anchor in human.xml
    return site


class HumanReplay:
    """One env's `_control_human` state machine (human_env.py:1685-1767): which clip and frame a control tick shows."""

    def __init__(self, tree, clips, clip_ids, human_pos_offset, human_rot_offset_wxyz, step_length):
        self.tree, self.clips, self.clip_ids = tree, clips, list(clip_ids)
        self.human_pos_offset, self.human_rot_offset = list(human_pos_offset), list(human_rot_offset_wxyz)
        self.step_length = step_length  # simulation_step_freq / human_animation_freq (human_env.py:1462-1465)
        self.index, self.animation_start_time, self.animation_time = 0, 0, -1

    def control_human(self, low_level_time):
        control_time = math.floor(low_level_time / self.step_length)
        self.animation_time = control_time - self.animation_start_time
        anim, info = self.clips[self.clip_ids[self.index]]
        if self.animation_time > anim["Pelvis_pos_x"].shape[0] - 1:
            self.index = (self.index + 1) % len(self.clip_ids)
            self.animation_time = 0
            self.animation_start_time = control_time
            anim, info = self.clips[self.clip_ids[self.index]]
        pos, quat = root_pose(anim, info, self.animation_time, self.human_pos_offset, self.human_rot_offset)
        angles = {k: anim[k][self.animation_time] for k in anim if k[-2:] in ("_x", "_y", "_z") and not k.startswith("Pelvis")}
        return site_positions(self.tree, pos, quat, angles)
