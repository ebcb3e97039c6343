"""Builds libofx.so (the C-ABI CUDA library, include/ofx.h) in-tree for sm_100a.

    python -m occlusionfusion_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels with the repo snapshot
to the GPU box.  --fmad=false: the only fused multiply-adds are the explicit __fmaf_rn calls,
so the device arithmetic is bit-identical to the contract.
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB_DIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIB_DIR, "libofx.so")
SOURCES = ["ofx_volume.cu", "ofx_points.cu", "ofx_skin_cache.cu", "ofx_integrate.cu", "ofx_graph.cu", "ofx_graph_edges.cu", "ofx_graph_host.cu", "ofx_mc.cu"]
HEADERS = ["ofx_mc_tables.h", "ofx_math.cuh", "ofx_volume.cuh", "ofx_brick.cuh", "ofx_fused.cuh", "ofx_scan.cuh", "ofx_common.h", os.path.join("..", "..", "include", "ofx.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "--fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-Xptxas", "-v"]
# -ffp-contract=off: the HOST code of the .cu files (ofx_graph_host.cu: geodesic path lengths and edge weights that must equal
# the reference's C++) may not be contracted into FMAs either -- GNU mode does that by default on aarch64 hosts.


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, out=None, defines=()):
    """out / defines: build an experimental variant (tools/ab_bench.sh) next to the product library."""
    if out is None and not force and not _stale():
        return LIB
    lib_path = out or LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    nvcc = _nvcc()
    env = dict(os.environ)
    if os.path.exists("/usr/bin/g++"):
        ccbin = ["-ccbin", "/usr/bin/g++"]
    else:
        ccbin = []
    objs = []
    logs = []
    for src in SOURCES:
        obj = os.path.join(LIB_DIR, src.replace(".cu", ".o" if out is None else "." + os.path.basename(out) + ".o"))
        cmd = [nvcc, *ccbin, *NVCC_FLAGS, *["-D" + d for d in defines], "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True, env=env)
        logs.append(f"$ {' '.join(cmd)}\n{r.stdout}{r.stderr}")
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + logs[-1])
        objs.append(obj)
    cmd = [nvcc, *ccbin, "-shared", "-o", lib_path, *objs]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    with open(os.path.join(LIB_DIR, "build.log" if out is None else os.path.basename(out) + ".log"), "w") as f:
        f.write("\n".join(logs))
    if verbose:
        print("\n".join(logs))
    return lib_path


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--out=")]
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, out=outs[0] if outs else None, defines=defs))
