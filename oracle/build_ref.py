#!/usr/bin/env python
"""Recipe for oracle/_ref/: a run-time copy of the reference's OWN implementation of the path, so that its CPU
torch code can be timed on the GPU box's host cores (where /root/reference does not exist).

    python oracle/build_ref.py        (also run by __graft_entry__.build() whenever /root/reference is present)

It copies /root/reference/models/utils.py byte for byte to oracle/_ref/reference_models_utils.py.  oracle/_ref/ is
listed in .gitignore (the reference's sources never enter this repository's history) but not in .gpurunignore, so
the copy travels to the GPU box next to the built .so files.  Test infrastructure only: ``oracle/ref_loader.py``
imports it for ``bench.py``'s CPU-baseline legs and for golden generation, never the product."""
import hashlib
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/models/utils.py"
DST_DIR = os.path.join(HERE, "_ref")
DST = os.path.join(DST_DIR, "reference_models_utils.py")


def build() -> str:
    """Returns the path of the copy ('' if neither the reference nor an earlier copy is available)."""
    if os.path.exists(SRC):
        os.makedirs(DST_DIR, exist_ok=True)
        if not os.path.exists(DST) or open(SRC, "rb").read() != open(DST, "rb").read():
            shutil.copyfile(SRC, DST)
        with open(os.path.join(DST_DIR, "SOURCE.txt"), "w") as fh:
            fh.write(f"{SRC}\nsha256 {hashlib.sha256(open(DST, 'rb').read()).hexdigest()}\n"
                     "verbatim copy made by oracle/build_ref.py; git-ignored, never committed\n")
    return DST if os.path.exists(DST) else ""


if __name__ == "__main__":
    print(build() or "reference not available")
