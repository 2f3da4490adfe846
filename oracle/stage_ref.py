"""Stage the UNMODIFIED reference package under the git-ignored oracle/_ref/ (TEST / BASELINE
INFRASTRUCTURE ONLY -- nothing of the product path reads it).

The reference is pure Python: there is nothing to compile.  This recipe copies
/root/reference/src/python_motion_planning (as it lies, no edits) to oracle/_ref/src/ so that the
CPU-baseline leg of bench.py can time the reference's own planners on the GPU box's host cores,
where /root/reference does not exist (oracle/_ref travels with gpurun like the built .so files; it
is listed in .gitignore, so no reference source ever enters the history).

    python oracle/stage_ref.py        # in the build container; __graft_entry__.build() calls it
"""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/src/python_motion_planning"
DST = os.path.join(HERE, "_ref", "src", "python_motion_planning")


def stage(verbose=False):
    if not os.path.isdir(SRC):
        return False                      # e.g. on the GPU box: use what was staged before
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    if verbose:
        n = sum(len(f) for _, _, f in os.walk(DST))
        print(f"staged {n} files of the reference under {DST}")
    return True


if __name__ == "__main__":
    stage(verbose=True)
