"""TEST/BUILD TOOLING (container only).  Condenses the reference's hand URDFs + ROBOT_CONFIG dicts into the
compact kinematic-tree JSON the product ships (compliancedex_b200/data/hands/<hand>.json).

Only the numbers the FK path reads are kept (urdf_utils.py:28-83: joint type / parent / child /
origin xyz+rpy / axis, in URDF *link order*; leap_hand_config.py:3-27, allegro_hand_config.py:3-27:
tip link names, tip offsets, ref_q, collision tables).  Run:  python -m oracle.make_hand_data
"""
import json
import os
import xml.etree.ElementTree as ET

from oracle import ref_harness as rh

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "compliancedex_b200", "data", "hands")


def urdf_to_tree(path):
    root = ET.parse(path).getroot()
    links = [l.get("name") for l in root.findall("link")]
    by_child = {}
    for j in root.findall("joint"):
        by_child[j.find("child").get("link")] = j
    bodies = []
    for i, name in enumerate(links):
        if i == 0:
            bodies.append(dict(name=name, parent=-1, type="fixed", xyz=[0.0, 0.0, 0.0], rpy=[0.0, 0.0, 0.0],
                               axis=[0.0, 0.0, 0.0], joint="base_joint"))
            continue
        j = by_child[name]
        o = j.find("origin")
        xyz = [float(v) for v in (o.get("xyz") if o is not None and o.get("xyz") else "0 0 0").split()]
        rpy = [float(v) for v in (o.get("rpy") if o is not None and o.get("rpy") else "0 0 0").split()]
        ax = j.find("axis")
        axis = [float(v) for v in ax.get("xyz").split()] if (ax is not None and j.get("type") != "fixed") else [0.0, 0.0, 0.0]
        bodies.append(dict(name=name, parent=links.index(j.find("parent").get("link")), type=j.get("type"),
                           xyz=xyz, rpy=rpy, axis=axis, joint=j.get("name")))
    return bodies


def main():
    os.makedirs(OUT, exist_ok=True)
    for hand in ["leap", "allegro"]:
        cfg = rh.hand_config(hand)
        doc = dict(hand=hand,
                   source="derived from the reference URDF + ROBOT_CONFIG (numbers only), see oracle/make_hand_data.py",
                   bodies=urdf_to_tree(rh.hand_urdf(hand)),
                   ee_link_name=list(cfg["ee_link_name"]),
                   ee_link_offset=cfg["ee_link_offset"].tolist(),
                   ref_q=cfg["ref_q"].tolist(),
                   collision_links=list(cfg["collision_links"]),
                   collision_offsets=cfg["collision_offsets"].tolist(),
                   collision_pairs=[list(map(int, p)) for p in cfg["collision_pairs"]])
        with open(os.path.join(OUT, f"{hand}.json"), "w") as f:
            json.dump(doc, f, indent=1)
        print("wrote", hand, len(doc["bodies"]), "bodies")


if __name__ == "__main__":
    main()
