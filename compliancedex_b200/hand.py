"""Hand description for the fused pregrasp kernels: the kinematic tree the reference builds from a URDF
(urdf_utils.py:28-83, robot_model.py:115-127) flattened into the arrays `cdx_hand_load` takes.

The tree comes either from the compact JSON shipped under data/hands/ (numbers extracted from the reference's
URDFs + ROBOT_CONFIG dicts) or from a URDF file given by path (stdlib XML; same link-order rule as the
reference: dof i drives the i-th non-fixed link in URDF *link* order).
"""
import json
import os
import xml.etree.ElementTree as ET

import numpy as np
import torch


def _rot32(axis, ang):
    """x_rot / y_rot / z_rot of spatial_vector_algebra.py:14-53 for one angle, float32 torch arithmetic."""
    ang = torch.as_tensor(ang, dtype=torch.float32)
    c, s = torch.cos(ang), torch.sin(ang)
    R = torch.zeros(3, 3, dtype=torch.float32)
    u, v = (axis + 1) % 3, (axis + 2) % 3
    R[axis, axis] = 1.0
    R[u, u] = c
    R[u, v] = -s
    R[v, u] = s
    R[v, v] = c
    return R


def bodies_from_urdf(path):
    """[{name,parent,type,xyz,rpy,axis}] in URDF link order (the order robot_model.py:115-127 walks)."""
    root = ET.parse(path).getroot()
    links = [l.get("name") for l in root.findall("link")]
    by_child = {j.find("child").get("link"): j for j in root.findall("joint")}
    bodies = []
    for i, name in enumerate(links):
        if i == 0:
            bodies.append(dict(name=name, parent=-1, type="fixed", xyz=[0.0] * 3, rpy=[0.0] * 3, axis=[0.0] * 3))
            continue
        j = by_child[name]
        o = j.find("origin")
        xyz = [float(v) for v in (o.get("xyz") if o is not None and o.get("xyz") else "0 0 0").split()]
        rpy = [float(v) for v in (o.get("rpy") if o is not None and o.get("rpy") else "0 0 0").split()]
        ax = j.find("axis")
        axis = [float(v) for v in ax.get("xyz").split()] if (ax is not None and j.get("type") != "fixed") else [0.0] * 3
        bodies.append(dict(name=name, parent=links.index(j.find("parent").get("link")), type=j.get("type"),
                           xyz=xyz, rpy=rpy, axis=axis))
    return bodies


class HandTree:
    """Flat arrays for cdx_hand_load (include/cdx.h)."""

    def __init__(self, bodies, ee_link_names, ee_link_offsets):
        self.names = [b["name"] for b in bodies]
        n = len(bodies)
        self.n_bodies = n
        self.parent = np.array([b["parent"] for b in bodies], dtype=np.int32)
        self.dof = np.full(n, -1, dtype=np.int32)
        self.axis = np.zeros(n, dtype=np.int32)
        self.sign = np.ones(n, dtype=np.float32)
        self.rot = np.zeros((n, 9), dtype=np.float32)
        self.xyz = np.zeros((n, 3), dtype=np.float32)
        nd = 0
        for i, b in enumerate(bodies):
            rpy = torch.tensor(b["rpy"], dtype=torch.float32)                       # urdf_utils.py:48-50
            self.xyz[i] = torch.tensor(b["xyz"], dtype=torch.float32).numpy()       # urdf_utils.py:51-53
            fixed = (_rot32(2, rpy[2]) @ _rot32(1, rpy[1])) @ _rot32(0, rpy[0])     # rigid_body.py:142
            self.rot[i] = fixed.reshape(-1).numpy()
            if i > 0 and b["type"] != "fixed":
                self.dof[i] = nd
                nd += 1
                ax = np.asarray(b["axis"], dtype=np.float32)
                a = 0 if abs(ax[0]) == 1 else (1 if abs(ax[1]) == 1 else 2)       # rigid_body.py:149-154
                self.axis[i] = a
                self.sign[i] = np.sign(ax[a])
            if i > 0 and not (0 <= b["parent"] < i):
                raise ValueError("kinematic tree must list parents before children (URDF link order)")
        self.n_dofs = nd
        if len(ee_link_names) != 4:
            raise ValueError("the pregrasp path is written for 4 fingertips (optimize_pregrasp.py:623-625)")
        self.tip_body = np.array([self.names.index(nm) for nm in ee_link_names], dtype=np.int32)
        self.tip_offset = np.asarray(ee_link_offsets, dtype=np.float32).reshape(4, 3)


def load_hand(hand_or_path):
    """'leap' / 'allegro' (shipped JSON) or a path to such a JSON -> dict (bodies, ee_link_name, ee_link_offset,
    ref_q, collision tables)."""
    path = hand_or_path
    if not os.path.isfile(str(path)):
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "hands", f"{hand_or_path}.json")
    with open(path) as f:
        return json.load(f)


def tree_for(robot_urdf, ee_link_names, ee_link_offsets):
    """robot_urdf: a URDF path (as the reference takes), a shipped hand name, a hand JSON path or a loaded dict."""
    if isinstance(robot_urdf, dict):
        bodies = robot_urdf["bodies"]
    elif str(robot_urdf).endswith((".urdf", ".xml")):
        bodies = bodies_from_urdf(robot_urdf)
    else:
        bodies = load_hand(robot_urdf)["bodies"]
    return HandTree(bodies, list(ee_link_names), ee_link_offsets)
