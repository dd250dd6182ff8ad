"""In-tree build of libscg_b200.so for sm_100a (nvcc cross-compiles without a GPU).

    python -m safe_control_gym_b200.build [--force] [-v]

One translation unit per (env kind, precision) plus the host ABI, compiled in parallel and linked
into one shared library.  The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
OUT = os.path.join(HERE, "libscg_b200.so")
STAMP = OUT + ".stamp"
NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC"] + os.environ.get("SCG_NVCC_EXTRA", "").split()   # e.g. -DSCG_NO_PAIRS=1 for an A/B build


def _headers():
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cuh", ".h"))]
    files.append(os.path.join(os.path.dirname(HERE), "include", "scg_b200.h"))
    return files


def _units():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]


def _digest(paths):
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for f in paths:
        h.update(open(f, "rb").read())
    return h.hexdigest()


def build(force=False, verbose=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ, exist_ok=True)
    heads = _headers()
    units = _units()
    total = _digest(heads + units)
    if not force and os.path.exists(OUT) and os.path.exists(STAMP) and open(STAMP).read().strip() == total:
        return OUT

    def compile_one(src):
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        stamp = obj + ".stamp"
        dig = _digest(heads + [src])
        if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read().strip() == dig:
            return obj, ""
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {os.path.basename(src)}:\n" + r.stderr[-6000:])
        with open(stamp, "w") as f:
            f.write(dig)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(len(units), os.cpu_count() or 4)) as ex:
        results = list(ex.map(compile_one, units))
    if verbose:
        for _, err in results:
            sys.stderr.write(err)
    r = subprocess.run([nvcc, "-shared", "-o", OUT] + [o for o, _ in results], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stderr[-4000:])
    with open(STAMP, "w") as f:
        f.write(total)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
