"""Snapshots the reference's own feature code into oracle/_ref/ -- TEST INFRASTRUCTURE ONLY.

The GPU box has no /root/reference, and the reference's CPU path for this hot path is pure Python
(AlphaGo/go.py + AlphaGo/preprocessing/preprocessing.py), so "compiling the reference from its own
sources into oracle/_ref/" is a verbatim copy of those files (plus the two empty package markers).
oracle/_ref/ is git-ignored (reference sources never enter this repository's history) but travels to
the GPU box with the snapshot, where bench.py's cpu_baseline leg times Preprocess.state_to_tensor
(preprocessing.py:266-294) -- the real thing -- beside the C port.  Nothing under rocalphago_b200/
ever imports it.

    python -m oracle.make_ref_snapshot          # run where /root/reference exists (the build container)
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
FILES = ["AlphaGo/__init__.py", "AlphaGo/go.py", "AlphaGo/preprocessing/__init__.py",
         "AlphaGo/preprocessing/preprocessing.py"]


def snapshot(reference_root="/root/reference"):
    if not os.path.isdir(os.path.join(reference_root, "AlphaGo")):
        return False
    for rel in FILES:
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(reference_root, rel), dst)
    with open(os.path.join(DEST, "README"), "w") as f:
        f.write("Verbatim copies of hiropppe/RocAlphaGo files made by oracle/make_ref_snapshot.py; git-ignored.\n")
    return True


def available():
    return all(os.path.exists(os.path.join(DEST, rel)) for rel in FILES)


if __name__ == "__main__":
    ok = snapshot(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("snapshot written to %s" % DEST if ok else "reference tree not found; nothing written")
