"""Builds tests/c_harness/abi_harness (plain C11 against include/bsx_b200.h + libbsx_b200.so) with gcc."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "c_harness", "abi_harness.c")
EXE = os.path.join(ROOT, "tests", "c_harness", "abi_harness")
LIBDIR = os.path.join(ROOT, "0-kno-blobstreamx_b200", "csrc")


def build():
    lib = os.path.join(LIBDIR, "libbsx_b200.so")
    if not os.path.exists(lib):
        raise RuntimeError("libbsx_b200.so is not built (python __graft_entry__.py)")
    if os.path.exists(EXE) and os.path.getmtime(EXE) > max(os.path.getmtime(SRC), os.path.getmtime(lib)):
        return EXE
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-O2", "-I" + os.path.join(ROOT, "include"), "-o", EXE, SRC,
           "-L" + LIBDIR, "-lbsx_b200", "-Wl,-rpath," + LIBDIR]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gcc failed:\n" + r.stderr)
    return EXE
