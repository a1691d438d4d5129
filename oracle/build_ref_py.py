"""TEST / BASELINE TOOLING: makes the reference's own Python for this path available on the GPU box.

/root/reference exists only in the build container.  This recipe copies the UNMODIFIED files the hot path consists of
(gpis.py, optimize_pregrasp.py, math_utils.py, the vendored differentiable_robot_model package, the two hand configs +
URDFs, the shipped GPIS states) into git-ignored oracle/_ref/reference_py/ - which, like the compiled torchsdf_ref.so,
travels to the GPU box with the snapshot but never enters the repository's history.  oracle/ref_harness.py then imports
the reference from there, so `bench.py --impl reference` and the cpu_baseline leg time the reference's own code
(`kind: "reference"`) on the box's host cores.  Nothing in compliancedex_b200/ ever touches it.

    python -m oracle.build_ref_py        (also run by __graft_entry__.build() when /root/reference is present)
"""
import glob
import os
import shutil

SRC = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_py")
ROBOTS = "pybullet_robot/src/pybullet_robot/robots"
FILES = ["gpis.py", "optimize_pregrasp.py", "math_utils.py",
         f"{ROBOTS}/leap_hand/leap_hand_config.py", f"{ROBOTS}/leap_hand/assets/leap_hand/robot.urdf",
         f"{ROBOTS}/allegro_hand/allegro_hand_config.py", f"{ROBOTS}/allegro_hand/models/allegro_hand_description_left.urdf",
         # robot_model.py:21-23 imports this data package only for its path; none of its 61 MB of robot meshes is needed
         "thirdparty/differentiable-robot-model/diff_robot_data/__init__.py"]
GLOBS = ["thirdparty/differentiable-robot-model/differentiable_robot_model/*.py", "gpis_states/*_state.npz"]


def build():
    if not os.path.isfile(os.path.join(SRC, "gpis.py")):
        return False
    rels = list(FILES)
    for g in GLOBS:
        rels += [os.path.relpath(p, SRC) for p in sorted(glob.glob(os.path.join(SRC, g)))]
    for rel in rels:
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), dst)
    with open(os.path.join(DST, "PROVENANCE.txt"), "w") as f:
        f.write("unmodified copies of %d files of %s made by oracle/build_ref_py.py; git-ignored, test/baseline use only\n" % (len(rels), SRC))
    return True


if __name__ == "__main__":
    print("reference_py built" if build() else "no /root/reference here: nothing copied")
