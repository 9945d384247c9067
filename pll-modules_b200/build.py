"""Build recipe for libpll_b200.so (product) and the oracle-side test infrastructure.

Everything is compiled IN-TREE with explicit nvcc / gcc command lines (no JIT cache), for
sm_100a only:  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3.
nvcc cross-compiles without a GPU, so this runs in the CPU-only build container too.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libpll_b200.so")
ORACLE = os.path.join(ROOT, "oracle")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function", "-Xptxas", "-v",
]

CU_SOURCES = ["device_ctx.cu", "api_core.cu", "k_pmatrix.cu", "k_partials.cu", "k_traversal_dna.cu", "k_traversal_dna4.cu", "trav_schedule.cpp", "k_traversal_aa.cu","k_likelihood.cu",
              "k_derivatives.cu", "k_newton.cu", "k_ancestral.cu", "binary_io.cpp", "host_model.cpp", "utree.cpp", "treeinfo.cpp", "optimize.cpp", "opt_algorithms.cpp", "opt_params.cpp", "algo_search.cpp", "algo_opt.cpp"]
C_SOURCES = ["state_maps.c"]


def _newer(target, deps):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _run(cmd, log=None):
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + res.stdout)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    return res.stdout


def build_library(force=False, verbose=False):
    """Compile the CUDA kernels + host code into pll-modules_b200/lib/libpll_b200.so."""
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, h) for h in os.listdir(CSRC) if h.endswith((".h", ".cuh"))]
    headers.append(os.path.join(ROOT, "include", "pllb200.h"))
    jobs = []
    objs = []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src + ".o")
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            jobs.append(([NVCC] + NVCC_FLAGS + ["-x", "cu", "-c", s, "-o", o], o + ".log"))
    for src in C_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src + ".o")
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            jobs.append((["gcc", "-O2", "-fPIC", "-c", s, "-o", o], o + ".log"))
    if jobs:
        with ThreadPoolExecutor(max_workers=8) as ex:
            outs = list(ex.map(lambda j: _run(j[0], j[1]), jobs))
        if verbose:
            for o in outs:
                print(o)
    if jobs or not os.path.exists(LIB):
        _run([NVCC, "-shared", "-o", LIB] + objs + ["-Xlinker", "--no-undefined", "-lpthread"])
    return LIB


def build_oracle():
    """Compile the oracle side: the C restatement, the same-source harness against the product
    library, and (only where /root/reference exists) the reference itself into oracle/_ref/."""
    out = {}
    restatement = os.path.join(ORACLE, "pll_oracle.c")
    if os.path.exists(restatement):
        so = os.path.join(ORACLE, "libpll_oracle.so")
        if not _newer(so, [restatement]):
            _run(["gcc", "-O2", "-fPIC", "-shared", "-fno-fast-math", restatement, "-o", so, "-lm"])
        out["restatement"] = so
    harness = os.path.join(ORACLE, "harness.c")
    if os.path.exists(harness) and os.path.exists(LIB):
        so = os.path.join(LIBDIR, "libharness_b200.so")
        if not _newer(so, [harness, LIB, os.path.join(ROOT, "include", "pllb200.h")]):
            _run(["gcc", "-O2", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include"), harness,
                  "-o", so, "-L" + LIBDIR, "-l:libpll_b200.so", "-Wl,-rpath,$ORIGIN", "-lm", "-lpthread"])
        out["harness_b200"] = so
    if os.path.isdir("/root/reference/libs/libpll/src"):
        ref_so = os.path.join(ORACLE, "_ref", "libpll_ref.so")
        h_so = os.path.join(ORACLE, "_ref", "libharness_ref.so")
        deps = [harness] if os.path.exists(harness) else []
        if not os.path.exists(ref_so) or (deps and not _newer(h_so, deps)):
            _run(["bash", os.path.join(ORACLE, "build_ref.sh")])
        out["reference"] = ref_so
    return out


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_oracle())
