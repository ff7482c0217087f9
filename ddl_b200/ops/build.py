"""Builds csrc/*.cu into ddl_b200/lib/libddl_b200.so for sm_100a with plain nvcc.

In-tree on purpose: the .so travels with the repo snapshot to the GPU box (a JIT cache under
~/.cache would not).  No torch headers are involved - the library is a plain C ABI loaded with
ctypes (ops/native.py), so a full rebuild takes seconds.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "csrc")
LIBDIR = os.path.join(ROOT, "ddl_b200", "lib")
LIB = os.path.join(LIBDIR, "libddl_b200.so")
LAB_LIB = os.path.join(LIBDIR, "libddl_b200_lab.so")     # same sources + -DDDL_TC_LAB: in-kernel traces / phase stamps / skip switches
                                                          # used by tools/tc_*.py, wgrad_trace.py, sync_phase.py (DDL_LAB=1)
OBJDIR = os.path.join(ROOT, "build", "obj")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found - cannot build the sm_100a library")
    return exe


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(lab: bool = False) -> str:
    h = hashlib.sha256()
    for path in sources() + [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cuh", ".h"))] \
            + [os.path.join(ROOT, "include", "ddl_b200.h")]:
        with open(path, "rb") as f:             # relative names: the stamp must survive a move of the tree
            h.update(os.path.relpath(path, ROOT).encode() + b"\0" + f.read())
    h.update(" ".join(NVCC_FLAGS + (["-DDDL_TC_LAB"] if lab else [])).encode())
    return h.hexdigest()


def needs_build(lab: bool = False) -> bool:
    lib = LAB_LIB if lab else LIB
    stamp = lib + ".sha256"
    if not (os.path.exists(lib) and os.path.exists(stamp)):
        return True
    return open(stamp).read().strip() != _digest(lab)


def build(force: bool = False, verbose: bool = False, lab: bool = False) -> str:
    LIB = LAB_LIB if lab else globals()["LIB"]
    if not force and not needs_build(lab):
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    exe = nvcc()
    logs = {}

    def compile_one(src):
        obj = os.path.join(OBJDIR, "%s%s.%d.o" % (os.path.basename(src)[:-3], ".lab" if lab else "", os.getpid()))
        cmd = [exe] + NVCC_FLAGS + (["-DDDL_TC_LAB"] if lab else []) + ["-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        logs[src] = r.stdout + r.stderr
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, logs[src]))
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(sources()))) as ex:
        objs = list(ex.map(compile_one, sources()))
    tmp = "%s.tmp.%d" % (LIB, os.getpid())        # link aside, then rename: concurrent ranks never see a torn .so
    link = [exe, "-shared", "-o", tmp] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    os.replace(tmp, LIB)
    for o in objs:
        try:
            os.remove(o)
        except OSError:
            pass
    with open(LIB + ".sha256.tmp.%d" % os.getpid(), "w") as f:
        f.write(_digest(lab))
    os.replace(LIB + ".sha256.tmp.%d" % os.getpid(), LIB + ".sha256")
    with open(os.path.join(ROOT, "build", "ptxas_lab.log" if lab else "ptxas.log"), "w") as f:
        for src, log in logs.items():
            f.write("==== %s\n%s\n" % (os.path.relpath(src, ROOT), log))
    if verbose:
        for src, log in logs.items():
            print("==== %s\n%s" % (os.path.relpath(src, ROOT), log))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, lab="--lab" in sys.argv))
    if "--sass" in sys.argv or "-v" in sys.argv:
        # per-kernel SASS evidence (profiles/sass/*.txt, profiles/sass_evidence.md), regenerated from the library just built
        subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "sass_extract.py")])
