"""Builds orb_slam_tracking_b200/liborbx.so (sm_100a only) in-tree with nvcc.

Parity-critical flags: -fmad=false (the reference's x86-64 baseline has no FMA; every float and
double expression must round after each multiply and add), IEEE division and square root,
no flush-to-zero.  Run as `python -m orb_slam_tracking_b200.build [--force]`.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OBJ = PKG / "build"
LIB = PKG / "liborbx.so"
SOURCES = ["pyramid.cu", "blur.cu", "fast.cu", "quadtree.cu", "retain_old.cu", "describe.cu", "glue.cu", "match.cu", "match_bf.cu", "score.cu", "tables.cu",
           "orbx_api.cu", "orbx_run.cu", "orbx_stream.cu", "no_alloc.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = (os.environ.get("ORBX_NVCC_EXTRA", "").split()) + ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
         "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-Xcompiler", "-fPIC,-O2,-Wall", "-Xptxas", "-v"]


def _newest_header():
    hs = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inc")) + [PKG.parent / "include" / "orbx.h"]
    return max(h.stat().st_mtime for h in hs)


def build(force=False, verbose=False):
    OBJ.mkdir(exist_ok=True)
    hdr = _newest_header()
    jobs = []
    for s in SOURCES:
        src, obj = CSRC / s, OBJ / (s + ".o")
        if force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr):
            jobs.append((src, obj))

    def cc(job):
        src, obj = job
        r = subprocess.run([NVCC, *FLAGS, "-c", str(src), "-o", str(obj)], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src.name, r.stderr))
        (OBJ / (src.name + ".ptxas.log")).write_text(r.stderr)
        if verbose:
            sys.stderr.write(r.stderr)

    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(cc, jobs))
    objs = [str(OBJ / (s + ".o")) for s in SOURCES]
    if jobs or not LIB.exists():
        r = subprocess.run([NVCC, "-shared", "-o", str(LIB), *objs, "-cudart", "static", "-Xlinker",
                            "--no-undefined"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
