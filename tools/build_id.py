"""Digest of the CUDA sources of librtb.so: csrc/*.cu, csrc/*.cuh (minus the generated build_id.h), csrc/Makefile and
include/rtb.h.  csrc/Makefile embeds it (rtb_build_id()); __graft_entry__ and tests/test_cabi.py compare the
library's digest with the tree's, so a stale prebuilt .so is caught.  usage: build_id.py [--header]"""
import glob
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def source_files():
    csrc = os.path.join(ROOT, "raytracing_b200", "csrc")
    files = sorted(glob.glob(os.path.join(csrc, "*.cu")) + glob.glob(os.path.join(csrc, "*.cuh")))
    files = [f for f in files if os.path.basename(f) != "build_id.h"]
    return files + [os.path.join(csrc, "Makefile"), os.path.join(ROOT, "include", "rtb.h")]


def digest():
    h = hashlib.sha256()
    for f in source_files():
        h.update(os.path.basename(f).encode() + b"\0")
        h.update(open(f, "rb").read())
    return h.hexdigest()[:32]


if __name__ == "__main__":
    d = digest()
    print('#define RTB_BUILD_ID "%s"' % d if "--header" in sys.argv else d)
