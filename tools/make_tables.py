"""Extract the numeric tables the hot path consumes from the reference's asset files.

Run in the build container only (needs /root/reference).  Output:
``fly_bproject_b200/data/nmf_tables.json`` — per asset: revolute DOF names in
URDF file order (Isaac Gym's real DOF order is not knowable offline, SURVEY.md
§8a), their limits (swapped if inverted, reference fly.py:258-264), link names
in file order, the local body indices of the abdomen / leg-tip links
(reference fly.py:299-300) and the default pose in radians (reference
fly.py:61-65, assets/pose_default.yaml, degrees).
"""
import json
import math
import os
import re
import sys

import yaml

REF = os.environ.get("FLYB200_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(os.path.dirname(__file__), "..", "fly_bproject_b200", "data", "nmf_tables.json")

ABDOMEN = ["A1A2", "A3", "A4", "A5", "A6"]                      # reference fly.py:299
LEG_TIPS = ["RFTarsus5", "LFTarsus5", "RMTarsus5", "LMTarsus5", "RHTarsus5", "LHTarsus5"]  # fly.py:300
ACTUATED = [  # reference fly.py:23-25
    "joint_LHCoxa_roll", "joint_RHCoxa_roll", "joint_LHFemur", "joint_RHFemur", "joint_LHTibia", "joint_RHTibia",
    "joint_LMCoxa_roll", "joint_RMCoxa_roll", "joint_LMFemur", "joint_RMFemur", "joint_LMTibia", "joint_RMTibia",
    "joint_LFCoxa", "joint_RFCoxa", "joint_LFFemur", "joint_RFFemur", "joint_LFTibia", "joint_RFTibia"]


def parse_urdf(path):
    text = open(path).read()
    links = re.findall(r'<link\s+name="([^"]+)"', text)
    dofs = []
    for m in re.finditer(r'<joint\s+name="([^"]+)"\s+type="([^"]+)">(.*?)</joint>', text, re.S):
        name, jtype, body = m.groups()
        if jtype != "revolute":
            continue
        lim = re.search(r'<limit[^>]*lower="([^"]+)"[^>]*upper="([^"]+)"', body)
        lo, hi = float(lim.group(1)), float(lim.group(2))
        if lo > hi:
            lo, hi = hi, lo
        dofs.append((name, lo, hi))
    return links, dofs


def main():
    pose_deg = yaml.safe_load(open(os.path.join(REF, "assets", "pose_default.yaml")))["joints"]
    out = {"actuated_names": ACTUATED, "abdomen_names": ABDOMEN, "leg_tip_names": LEG_TIPS, "assets": {}}
    for key, fname in (("limited_18", "nmf_no_limits_limited_Dofs.urdf"), ("full_42", "nmf_no_limits.urdf")):
        links, dofs = parse_urdf(os.path.join(REF, "assets", fname))
        names = [d[0] for d in dofs]
        out["assets"][key] = {
            "urdf": "assets/" + fname,
            "num_dof": len(dofs),
            "num_bodies": len(links),
            "dof_names": names,
            "lower": [d[1] for d in dofs],
            "upper": [d[2] for d in dofs],
            "initial_pos": [math.radians(pose_deg.get(n, 0.0)) for n in names],
            "action_index_18": sorted(names.index(n) for n in ACTUATED),
            "abdomen_local": sorted(links.index(n) for n in ABDOMEN),
            "leg_tip_local": sorted(links.index(n) for n in LEG_TIPS),
        }
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    for k, v in out["assets"].items():
        print(k, v["num_dof"], v["num_bodies"], v["abdomen_local"], v["leg_tip_local"], v["action_index_18"])


if __name__ == "__main__":
    sys.exit(main())
