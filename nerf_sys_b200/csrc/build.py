"""Build libacnerf_sm100a.so in-tree with nvcc (sm_100a only).

    python nerf_sys_b200/csrc/build.py [--force] [--verbose]

Per-file flags: the geometry / encoder files are compiled with -fmad=false so every
multiply and add rounds separately, exactly like the reference's elementwise torch
ops (the floor() of a scaled coordinate decides a hash-table row); the MLP /
compositing files keep FMA contraction.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIB = os.path.join(PKG, "libacnerf_sm100a.so")
OBJ = os.path.join(HERE, "build")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "--expt-relaxed-constexpr", "-Xptxas", "-v"]
SOURCES = {
    "api.cu": [],
    "rays.cu": ["-fmad=false"],
    "routing.cu": ["-fmad=false"],
    "hashgrid.cu": ["-fmad=false"],
    "wgrad.cu": [],
    "ngp.cu": [],
    "ngp_tc.cu": [],
    "dense_tc.cu": [],
    "render.cu": [],
    "optim.cu": [],
    "a2a.cu": ["-fmad=false"],
    "occ.cu": ["-fmad=false"],
    "episode.cu": ["-fmad=false"],
    "voronoi.cu": ["-fmad=false"],
}


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    raise RuntimeError("nvcc not found")


def _digest(path, extra):
    h = hashlib.sha256()
    for f in sorted(os.listdir(HERE)):
        if f.endswith((".cuh", ".h")):
            h.update(open(os.path.join(HERE, f), "rb").read())
    h.update(open(os.path.join(os.path.dirname(PKG), "include", "acnerf.h"), "rb").read())
    h.update(open(path, "rb").read())
    h.update(" ".join(extra).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    objs, jobs = [], []
    for src, extra in SOURCES.items():
        path = os.path.join(HERE, src)
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        stamp = obj + ".sha"
        dg = _digest(path, extra)
        objs.append(obj)
        if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dg:
            continue
        jobs.append((src, [nvcc, *ARCH, *COMMON, *extra, "-c", path, "-o", obj], stamp, dg))

    def run(job):
        src, cmd, stamp, dg = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        with open(stamp, "w") as f:
            f.write(dg)
        with open(os.path.join(OBJ, src + ".ptxas.txt"), "w") as f:
            f.write(r.stderr)
        return src, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src, log in ex.map(run, jobs):
            if verbose:
                print(f"== {src}\n{log}")
    if jobs or not os.path.exists(LIB) or force:
        cmd = [nvcc, *ARCH, "-shared", "-o", LIB, *objs, "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
