"""Builds pysdm_b200/libpysdm_b200.so (hand-written CUDA for sm_100a + the C-ABI) in-tree
with nvcc; cross-compiles without a GPU.  `python -m pysdm_b200.build [--force]`.

Every translation unit is compiled to an object file under pysdm_b200/build/ (git-ignored) in
parallel and only when it is older than its sources, then linked into the shared library."""

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
SO = os.path.join(HERE, "libpysdm_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
NVCC_FLAGS = (
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # no implicit FMA contraction: results must match the reference's IEEE sequence;
    # fma() is written explicitly where the reference itself fuses
    "-fmad=false",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
)
LINK_FLAGS = ("-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static",
              "-Xcompiler", "-fPIC")


def nvcc_path():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.endswith(".cu")]
    deps.append(os.path.join(INCLUDE, "pysdm_b200.h"))
    return deps


def is_stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(d) > t for d in sources() + _headers())


def build_extension(force=False, verbose=False, out=None, defines=()):
    """`out`/`defines`: tuning variants (e.g. -DPYSDM_WARP_POS=768) built next to the default"""
    if out is None and not force and not is_stale():
        return SO
    variant = out is not None
    out = out or SO
    obj_dir = OBJ if not variant else OBJ + "_" + os.path.basename(out).replace(".", "_")
    os.makedirs(obj_dir, exist_ok=True)
    newest_header = max(os.path.getmtime(h) for h in _headers())
    nvcc = nvcc_path()

    def compile_one(src):
        obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
        if (not force and not variant and os.path.exists(obj)
                and os.path.getmtime(obj) > max(os.path.getmtime(src), newest_header)):
            return obj
        cmd = [nvcc, *NVCC_FLAGS, *defines, "-I", CSRC, "-I", INCLUDE, "-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd))
        subprocess.check_call(cmd)
        return obj

    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as pool:
        objects = list(pool.map(compile_one, sources()))
    subprocess.check_call([nvcc, *LINK_FLAGS, "-o", out, *objects])
    return out


if __name__ == "__main__":
    _out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
    _defs = tuple(a for a in sys.argv[1:] if a.startswith("-D"))
    print(build_extension(force="--force" in sys.argv, verbose="-v" in sys.argv, out=_out,
                          defines=_defs))
