"""Regenerates tests/golden/human_tree.json and tests/golden/cmu_clips.npz (run in the build container only).

/root/reference does not exist on the GPU box, so what the animation -> keypoints tests need from it is committed as two
small fixtures: the kinematic tree parsed from the reference's HRG/models/assets/human/human.xml (names, parents, anchors,
in HumanObject.joint_elements order), and the head of two of its recorded clips (CMU/62_01: 400 frames, CMU/62_03: 60
frames; all 73 channels, FP64) with their *_info.json content. No MuJoCo here: expected keypoints come from
oracle/human_fk_oracle.py at test time (parity unpinned, see its header).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from human_robot_gym_b200.human_animation import JOINT_ELEMENTS, load_clip, load_human_tree, reorder_tree  # noqa: E402

REF = os.environ.get("HRG_REFERENCE", "/root/reference")
HUMAN = os.path.join(REF, "human_robot_gym/models/assets/human")
CLIPS = [("CMU/62_01", 400), ("CMU/62_03", 60)]


def main():
    tree = reorder_tree(load_human_tree(os.path.join(HUMAN, "human.xml")), JOINT_ELEMENTS)
    with open(os.path.join(HERE, "human_tree.json"), "w") as f:
        json.dump(tree, f, indent=0)
    out = {}
    for i, (name, n) in enumerate(CLIPS):
        anim, info = load_clip(os.path.join(HUMAN, "animations/human-robot-animations", name + ".pkl"))
        keys = sorted(anim)
        out[f"keys{i}"] = np.array(keys)
        out[f"quat{i}"] = anim["Pelvis_quat"][:n]
        out[f"chan{i}"] = np.stack([anim[k][:n] for k in keys if k != "Pelvis_quat"], axis=1)
        out[f"info{i}"] = np.array(list(info["position_offset"]) + list(info["orientation_quat"]), dtype=np.float64)
        out[f"name{i}"] = np.array(name)
    np.savez_compressed(os.path.join(HERE, "cmu_clips.npz"), **out)
    print("wrote human_tree.json,", "cmu_clips.npz", os.path.getsize(os.path.join(HERE, "cmu_clips.npz")), "bytes")


if __name__ == "__main__":
    sys.exit(main())
