"""Builds libcgnn_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension).

The shared library is the product; it is built here (cross-compiled, no GPU needed) and travels
to the GPU box with the repository snapshot.  Translation units are compiled in parallel:
capi.cu (pairwise path host side, life cycle, probes), pairwise_inst.cu once per CTA size of the pairwise kernel
(512 / 256 / 128 threads), graph_api.cu (graph path host side) and graph_inst.cu
once per (padded column count, padded hidden width, noise mode) (-DCGNN_DP=8|16|20|22|24|32 -DCGNN_HP=20|32
-DCGNN_REGEN=0|1).
"""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libcgnn_b200.so")
GRAPH_DPS = (8, 16, 20, 22, 24, 32)
GRAPH_HPS = (20, 32)
PAIRWISE_BUILDS = ((512, 2), (256, 4), (128, 8))      # threads per CTA, CTAs per SM
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def _units():
    u = [("capi", "capi.cu", []), ("graph_api", "graph_api.cu", [])]
    u += [("pairwise_%d" % nt, "pairwise_inst.cu", ["-DCGNN_PW_NT=%d" % nt, "-DCGNN_PW_CTAS=%d" % ctas, "-DCGNN_PW_NS=pw%d" % nt,
                                                     "-DCGNN_PW_TAG=%d" % nt]) for nt, ctas in PAIRWISE_BUILDS]
    u += [("graph_dp%d_hp%d%s" % (d, hp, "_regen" if rg else ""), "graph_inst.cu",
           ["-DCGNN_DP=%d" % d, "-DCGNN_HP=%d" % hp, "-DCGNN_REGEN=%d" % rg])
          for rg in (0, 1) for d in GRAPH_DPS for hp in GRAPH_HPS]
    return u


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libcgnn_b200.so")
    return exe


def _deps_mtime():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "cgnn_b200.h"))
    return max(os.path.getmtime(d) for d in deps if os.path.exists(d))


def source_hash(kind):
    """Hash of the CUDA sources a kernel family is compiled from ('pairwise' | 'graph').  profiles/*.json written by
    tools/ncu_summary.py carry it, bench.py recomputes it: a profile of an older kernel is flagged, not trusted."""
    import hashlib
    files = {"pairwise": ("common.cuh", "pairwise.cuh"), "graph": ("common.cuh", "graph.cuh")}[kind]
    h = hashlib.sha256()
    for f in files:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def build_library(force=False, verbose=False):
    """Compile the CUDA sources into cgnn_b200/libcgnn_b200.so; returns its path."""
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _deps_mtime():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    logs = []

    def compile_one(unit):
        name, src, defs = unit
        obj = os.path.join(OBJ, name + ".o")
        cmd = [nvcc] + NVCC_FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s%s" % (src, r.stdout, r.stderr))
        logs.append("== %s\n%s" % (name, r.stderr))
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, _units()))
    r = subprocess.run([nvcc, "-shared", "-o", LIB] + objs, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    if verbose:
        print("\n".join(logs))
    return LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
