"""Stages the reference's Python modules where they can travel to the GPU box.  TEST INFRASTRUCTURE ONLY.

    python -m oracle.stage_ref          (also run by __graft_entry__.build() when /root/reference exists)

/root/reference exists only in the authoring container.  The CPU arm of bench.py
(`--impl reference`, `cpu_baseline`) times the reference's OWN training loop on the GPU box's host
cores, so its `modules/` tree is copied, unmodified, into oracle/_ref/modules.  oracle/_ref/ is
listed in .gitignore (no reference source enters the history) but not in .gpurunignore (it travels
with the snapshot, like the built libbqa.so).  oracle/reference_shims.py imports it from there when
/root/reference is absent.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("BQA_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")


def stage(verbose=True):
    src_modules = os.path.join(SRC, "modules")
    if not os.path.isdir(src_modules):
        if verbose:
            print(f"stage_ref: {src_modules} not found, nothing staged", file=sys.stderr)
        return False
    dst_modules = os.path.join(DST, "modules")
    if os.path.isdir(dst_modules):
        shutil.rmtree(dst_modules)
    os.makedirs(DST, exist_ok=True)
    shutil.copytree(src_modules, dst_modules, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    with open(os.path.join(DST, "STAGED_FROM"), "w") as f:
        f.write(f"unmodified copy of {src_modules} (test infrastructure; git-ignored)\n")
    if verbose:
        n = sum(len(fs) for _, _, fs in os.walk(dst_modules))
        print(f"stage_ref: {n} files -> {dst_modules}")
    return True


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
