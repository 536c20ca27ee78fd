"""In-tree build of liboflow_b200.so (nvcc, sm_100a only).

``python -m optical_flow_b200.build`` or ``build_library()``.  Objects are
compiled in parallel (one nvcc per .cu) and only when stale.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "liboflow_b200.so")
LIB_PROF = os.path.join(HERE, "liboflow_b200_prof.so")      # -DOF_PROFILE_HOOKS build (profiling scripts only)
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc():
    nv = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nv):
        raise RuntimeError("nvcc not found")
    return nv


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps_mtime():
    m = 0.0
    for root in (CSRC, INCLUDE):
        for f in os.listdir(root):
            if f.endswith((".cuh", ".h")):
                m = max(m, os.path.getmtime(os.path.join(root, f)))
    return m


def build_library(force: bool = False, verbose: bool = False, profile_hooks: bool = False, tag: str = "",
                  defines=()) -> str:
    """profile_hooks: build liboflow_b200_prof.so with -DOF_PROFILE_HOOKS (environment-variable tuning knobs and the
    "skip part of the kernel" debug paths; see csrc/common.cuh).  The shipped library has none of them.
    tag / defines: an A/B variant liboflow_b200_<tag>.so compiled with extra -D macros (tuning scripts only; selected
    at run time with OF_PROFILE_LIB=<tag>)."""
    build_dir = BUILD + ("_prof" if profile_hooks else "") + ("_" + tag if tag else "")
    lib_path = LIB_PROF if profile_hooks else LIB
    if tag:
        lib_path = os.path.join(HERE, "liboflow_b200_%s.so" % tag)
    extra = (["-DOF_PROFILE_HOOKS"] if profile_hooks else []) + ["-D" + d for d in defines]
    os.makedirs(build_dir, exist_ok=True)
    nvcc = _nvcc()
    hdr_m = _deps_mtime()
    jobs = []
    objs = []
    for src in _sources():
        s = os.path.join(CSRC, src)
        o = os.path.join(build_dir, src[:-3] + ".o")
        objs.append(o)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(s), hdr_m):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        cmd = [nvcc] + NVCC_FLAGS + extra + ["-I", INCLUDE, "-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = o[:-2] + ".ptxas.log"
        with open(log, "w") as f:
            f.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (s, r.stdout + r.stderr))
        return s

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for s in ex.map(compile_one, jobs):
                if verbose:
                    print("compiled", os.path.basename(s))
    if jobs or not os.path.exists(lib_path):
        cmd = [nvcc, "-shared", "-o", lib_path] + objs + ["-lcudart", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return lib_path


if __name__ == "__main__":
    _tag = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--tag=")), "")
    _defs = [a[2:] for a in sys.argv if a.startswith("-D")]
    print(build_library(force="--force" in sys.argv, verbose=True, profile_hooks="--profile-hooks" in sys.argv,
                        tag=_tag, defines=_defs))
