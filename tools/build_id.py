"""Digest of the CUDA sources of librtb.so: csrc/*.cu, csrc/*.cuh (minus the generated build_id.h), csrc/Makefile and
include/rtb.h.  csrc/Makefile embeds it (rtb_build_id()); __graft_entry__ and tests/test_cabi.py compare the
library's digest with the tree's, so a stale prebuilt .so is caught.  A second digest covers only the sources of the
trace kernels (rtb_kernel_id()).  usage: build_id.py [--header]"""
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


KERNEL_SOURCES = ("cell_tables.cuh", "fastmath.cuh", "field.cuh", "golden_table.cuh", "step.cuh", "step_fast.cuh",
                  "traj_stage.cuh", "trace_kernel.cuh", "trace_inst.cu")


def kernel_files():
    """The sources the trace kernels are compiled from (not the host side of the library): the key of the ncu captures
    in profiles/ncu_latest.json, so a host-only change does not orphan them."""
    csrc = os.path.join(ROOT, "raytracing_b200", "csrc")
    return [os.path.join(csrc, f) for f in KERNEL_SOURCES]


def _digest(files):
    h = hashlib.sha256()
    for f in files:
        h.update(os.path.basename(f).encode() + b"\0")
        h.update(open(f, "rb").read())
    return h.hexdigest()[:32]


def digest():
    return _digest(source_files())


def kernel_digest():
    return _digest(kernel_files())


if __name__ == "__main__":
    if "--header" in sys.argv:
        print('#define RTB_BUILD_ID "%s"\n#define RTB_KERNEL_ID "%s"' % (digest(), kernel_digest()))
    else:
        print(digest(), kernel_digest())
