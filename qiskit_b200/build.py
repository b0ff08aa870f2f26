"""Build recipe for libqsv_b200.so (nvcc, sm_100a, in-tree).

    python -m qiskit_b200.build            # rebuild the objects whose sources changed, relink
    python -m qiskit_b200.build --force
    python -m qiskit_b200.build -v         # ptxas -v (registers / spills per kernel)

Every translation unit is compiled to its own object (in parallel) under ``csrc/_obj`` and the
objects are linked into the shared library; a unit is recompiled when it or a header is newer
than its object.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libqsv_b200.so")
SOURCES = ["qsv_state.cu", "qsv_gate.cu", "qsv_expval.cu", "qsv_sample.cu", "qsv_dist.cu", "qsv_density.cu",
           "qsv_counts.cu", "qsv_graph.cu", "qsv_noalloc.cu"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
]


def _headers_mtime() -> float:
    paths = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    paths.append(os.path.join(HERE, "..", "include", "qsv.h"))
    return max(os.path.getmtime(p) for p in paths)


# kernels that run at the 128-register limit of two 256-thread CTAs per SM: a spill there costs 25 % on
# every benchmark stream (measured in round 2 when a macro sub-op pushed k_fused_pass over the limit)
NO_SPILL = ("k_fused_pass", "k_fused_swap_pass")


def _check_spills(log: str):
    lines = log.splitlines()
    for i, line in enumerate(lines):
        if "Function properties for" in line and any(k in line for k in NO_SPILL):
            props = lines[i + 1] if i + 1 < len(lines) else ""
            if "spill stores" in props and ", 0 bytes spill stores" not in props:
                raise RuntimeError(f"register spills in a kernel that must not spill: {line.strip()} -> {props.strip()}")


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ, exist_ok=True)
    hdr = _headers_mtime()
    sources = [f for f in SOURCES if os.path.exists(os.path.join(CSRC, f))]
    jobs = []
    for f in sources:
        src, obj = os.path.join(CSRC, f), os.path.join(OBJ, f[:-3] + ".o")
        stale = force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr)
        if stale:
            cmd = [nvcc, *NVCC_FLAGS, "-Xptxas", "-v", "-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        return cmd, subprocess.run(cmd, capture_output=True, text=True)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as pool:
            results = list(pool.map(run, jobs))
        for cmd, res in results:
            if res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
                raise RuntimeError("nvcc failed: " + " ".join(cmd))
            if verbose:
                sys.stderr.write(res.stdout + res.stderr)
            try:
                _check_spills(res.stdout + res.stderr)
            except RuntimeError:
                os.remove(cmd[cmd.index("-o") + 1])   # never link (or keep as up to date) an object that spills
                raise
    objs = [os.path.join(OBJ, f[:-3] + ".o") for f in sources]
    if jobs or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(o) for o in objs):
        cmd = [nvcc, "--shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a",
               *objs, "-o", LIB, "-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("linking libqsv_b200.so failed")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
