"""Builds pysdm_b200/libpysdm_b200.so (hand-written CUDA for sm_100a + the C-ABI) in-tree
with nvcc; cross-compiles without a GPU.  `python -m pysdm_b200.build [--force]`."""

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libpysdm_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
SOURCES = ("index.cu", "pairs.cu", "storage_ops.cu", "physics.cu", "breakup.cu", "transport.cu",
           "migration.cu", "seeding.cu", "fused.cu")
NVCC_FLAGS = (
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # no implicit FMA contraction: results must match the reference's IEEE sequence;
    # fma() is written explicitly where the reference itself fuses
    "-fmad=false",
    "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared", "-cudart", "static",
)


def nvcc_path():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def is_stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(INCLUDE, "pysdm_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build_extension(force=False, verbose=False, out=None, defines=()):
    """`out`/`defines`: tuning variants (e.g. -DPYSDM_WARP_POS=768) built next to the default"""
    if out is None and not force and not is_stale():
        return SO
    out = out or SO
    cmd = [nvcc_path(), *NVCC_FLAGS, *defines, "-I", CSRC, "-I", INCLUDE, "-o", out, *sources()]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    _out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
    _defs = tuple(a for a in sys.argv[1:] if a.startswith("-D"))
    print(build_extension(force="--force" in sys.argv, verbose="-v" in sys.argv, out=_out,
                          defines=_defs))
