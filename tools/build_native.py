#!/usr/bin/env python3
"""Build the in-tree native libraries.

* ``hetseq_9cme_b200/lib/libb200seq.so``      — every ``csrc/*.cu`` compiled for sm_100a only
  (``-gencode arch=compute_100a,code=sm_100a -lineinfo``), linked into one C-ABI shared object that the
  Python side loads with ctypes (see ``hetseq_9cme_b200/ops/native.py``).
* ``hetseq_9cme_b200/lib/libb200seq_host.so`` — ``csrc/host/*.cpp`` (batcher, shard index): no CUDA
  dependency, so the data layer works on machines without a driver.

The objects are compiled in parallel and cached by source mtime under ``build/``.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT = ROOT / "hetseq_9cme_b200" / "lib"
BUILD = ROOT / "build"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"),
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: Path, sources) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in sources)


def _run(cmd):
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(map(str, cmd)), proc.stdout, proc.stderr))
    return proc.stdout + proc.stderr


def build_cuda(verbose: bool = False, force: bool = False) -> Path:
    OUT.mkdir(parents=True, exist_ok=True)
    BUILD.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    headers = list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    sources = sorted(CSRC.glob("*.cu"))
    jobs = []
    objs = []
    for src in sources:
        obj = BUILD / (src.stem + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", str(src), "-o", str(obj)]
            jobs.append(cmd)
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            for out in ex.map(_run, jobs):
                if verbose and out.strip():
                    print(out)
    lib = OUT / "libb200seq.so"
    if jobs or _stale(lib, objs):
        _run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(lib)] + [str(o) for o in objs]
             + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    return lib


def build_host(force: bool = False) -> Path:
    OUT.mkdir(parents=True, exist_ok=True)
    sources = sorted((CSRC / "host").glob("*.cpp"))
    lib = OUT / "libb200seq_host.so"
    if sources and (force or _stale(lib, sources)):
        cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
        _run([cxx, "-O3", "-std=c++17", "-fPIC", "-shared", "-I", str(ROOT / "include"), "-o", str(lib)]
             + [str(s) for s in sources] + ["-lpthread"])
    return lib


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    verbose = "-v" in argv
    force = "-f" in argv
    print("host  :", build_host(force))
    print("cuda  :", build_cuda(verbose, force))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
