#!/usr/bin/env python
"""Recipe for `oracle/_ref/`: the reference is one Python file with no build step, so "building" it is copying
`/root/reference/server.py` - unmodified - to `oracle/_ref/server.py`.  `oracle/_ref/` is git-ignored (reference sources
never enter this repository's history) but travels with the working tree to the GPU box, where `bench.py --impl
reference` and the `cpu_baseline` leg run it on the host cores through `oracle/ref_harness.py`.
Run by `__graft_entry__.build()` wherever `/root/reference` exists; a no-op elsewhere."""
import os
import shutil
import sys

SRC = "/root/reference/server.py"
DST_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def build_ref() -> bool:
    if not os.path.exists(SRC):
        return os.path.exists(os.path.join(DST_DIR, "server.py"))
    os.makedirs(DST_DIR, exist_ok=True)
    shutil.copyfile(SRC, os.path.join(DST_DIR, "server.py"))
    return True


if __name__ == "__main__":
    print("oracle/_ref/server.py", "present" if build_ref() else "absent (no /root/reference here)")
    sys.exit(0)
