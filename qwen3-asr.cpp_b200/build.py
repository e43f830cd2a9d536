"""Builds libq3a.so (the product: CUDA kernels + C ABI) in-tree with nvcc for sm_100a.

    python qwen3-asr.cpp_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The library links cudart statically and resolves the one driver
entry point it needs (cuTensorMapEncodeTiled) at run time, so it loads on a box without a driver;
compute calls then fail with Q3A_ERR_CUDA.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libq3a.so")
OBJ = os.path.join(HERE, "build")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-O3", "--expt-relaxed-constexpr", "-Xptxas", "-v",
         "-DQ3A_BUILD"]


def sources() -> list[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def headers() -> list[str]:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "q3a.h"))
    return hs


def source_hash() -> str:
    """sha256 over the kernel sources (csrc/* and include/q3a.h, names + contents): ties a committed profile to the
    build it was taken from (profiles/roofline_traffic.json carries it; bench.py prints traffic: null on mismatch)."""
    import hashlib
    h = hashlib.sha256()
    for path in sorted(sources() + headers()):
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hs = headers()
    jobs = []
    objs = []
    for src in sources():
        obj = os.path.join(OBJ, os.path.basename(src) + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + hs + [os.path.abspath(__file__)]):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [NVCC, *FLAGS, "-x", "cu", "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, r in ex.map(compile_one, jobs):
                log = os.path.join(OBJ, os.path.basename(src) + ".log")
                with open(log, "w") as f:
                    f.write(r.stdout + r.stderr)
                if verbose or r.returncode != 0:
                    sys.stderr.write(r.stdout + r.stderr)
                if r.returncode != 0:
                    raise RuntimeError(f"nvcc failed on {src} (see {log})")
    if force or jobs or _stale(OUT, objs):
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT, *objs,
               "-cudart", "static", "-Xlinker", "--no-undefined", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    if "--hash" in sys.argv:
        print(source_hash())
        sys.exit(0)
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
