"""Stage the reference's own Python files for machines that have no /root/reference (TEST INFRASTRUCTURE ONLY).

The GPU box receives a snapshot of this repository and nothing else. Tests and the `--impl reference` bench arm
that need the UNMODIFIED reference modules / scripts there (models/*.py, main.py, main_dist.py, utils.py) read
them from `oracle/_ref/reference/`, which this recipe fills from the reference tree where it lies
(`/root/reference`, read-only). `oracle/_ref/` is git-ignored — reference sources never enter this repository's
history — but it is not gpurun-ignored, so it travels with the snapshot like the built libb2c.so.
`__graft_entry__.build()` runs this whenever /root/reference is present. Nothing under pytorch-cifar_b200/ reads
these files.
"""
import glob
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SOURCE = os.environ.get("B2C_REFERENCE_SRC", "/root/reference")
STAGED = os.path.join(HERE, "_ref", "reference")
TOP_LEVEL = ("main.py", "main_dist.py", "utils.py")


def stage(verbose=False):
    """Copy the reference's scripts and model files into oracle/_ref/reference. Returns the staged directory, or
    None when the reference tree is not on this machine (then whatever was staged earlier is left alone)."""
    if not os.path.isdir(os.path.join(SOURCE, "models")):
        return None
    os.makedirs(os.path.join(STAGED, "models"), exist_ok=True)
    files = [os.path.join(SOURCE, f) for f in TOP_LEVEL] + sorted(glob.glob(os.path.join(SOURCE, "models", "*.py")))
    for src in files:
        dst = os.path.join(STAGED, os.path.relpath(src, SOURCE))
        shutil.copyfile(src, dst)
        if verbose:
            print("staged", dst)
    return STAGED


def reference_dir():
    """Directory holding the reference's files on this machine: the real tree if present, else the staged copy,
    else None."""
    if os.path.isdir(os.path.join(SOURCE, "models")):
        return SOURCE
    if os.path.isdir(os.path.join(STAGED, "models")):
        return STAGED
    return None


if __name__ == "__main__":
    print(stage(verbose=True) or f"{SOURCE} not found; nothing staged")
