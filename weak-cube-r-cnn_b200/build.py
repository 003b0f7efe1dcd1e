"""Build libwcr_b200.so (hand-written CUDA for sm_100a) in-tree with nvcc.

    python weak-cube-r-cnn_b200/build.py [--force]
    python weak-cube-r-cnn_b200/build.py --prof      # + libwcr_b200_prof.so (phase cycle counters, diagnostics)

Flags that matter:
  -gencode arch=compute_100a,code=sm_100a   B200 only, no PTX fallback for others
  -fmad=false   no implicit FMA contraction: the only fused multiply-adds are the
                explicit fmaf()/fma.rn.f32x2 sites, which is what makes counts
                and indices bit-comparable with oracle/ (DESIGN.md "Rounding")
  -lineinfo     ncu source page maps to these files
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "wcr", "_lib")
OUT = os.path.join(OUT_DIR, "libwcr_b200.so")
# the first, unculled scoring kernel as an independent cross-check (tests only, never the product):
# the same sources with -DWCR_KEEP_BASELINE_KERNEL, where wcr_score_topk runs score_items_kernel
OUT_BASELINE = os.path.join(OUT_DIR, "libwcr_b200_baseline.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2",
    "--expt-relaxed-constexpr",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _stale() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(ROOT, "include", "*.h")) + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    if not force and not _stale() and os.path.exists(OUT_BASELINE) and \
            os.path.getmtime(OUT_BASELINE) >= os.path.getmtime(OUT):
        return OUT
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    procs = []
    base_obj = os.path.join(OUT_DIR, "score_baseline.o")
    procs.append(("score.cu (baseline)", subprocess.Popen(
        [nvcc, *NVCC_FLAGS, "-DWCR_KEEP_BASELINE_KERNEL", "-c", os.path.join(CSRC, "score.cu"), "-o", base_obj],
        stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src in sources():
        obj = os.path.join(OUT_DIR, os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(out.decode())
        if p.returncode != 0:
            failed = True
    if failed:
        raise RuntimeError("nvcc failed")
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
                           "-o", OUT, *objs, "-lcudart"])
    base_objs = [base_obj if os.path.basename(o) == "score.o" else o for o in objs]
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
                           "-o", OUT_BASELINE, *base_objs, "-lcudart"])
    return OUT


def build_prof() -> str:
    """libwcr_b200_prof.so: the product sources with -DWCR_PROFILE_PHASES (score_tiles_kernel adds up
    the SM cycles of its phases in the workspace header). Diagnostics only
    (tools/diag_score.py --lib=libwcr_b200_prof.so); never loaded by the package."""
    build()
    nvcc = os.environ.get("NVCC", "nvcc")
    out = os.path.join(OUT_DIR, "libwcr_b200_prof.so")
    obj = os.path.join(OUT_DIR, "score_prof.o")
    subprocess.check_call([nvcc, *NVCC_FLAGS, "-DWCR_PROFILE_PHASES", "-c", os.path.join(CSRC, "score.cu"), "-o", obj])
    objs = [obj if os.path.basename(s)[:-3] == "score" else os.path.join(OUT_DIR, os.path.basename(s)[:-3] + ".o")
            for s in sources()]
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out, *objs, "-lcudart"])
    return out


if __name__ == "__main__":
    if "--prof" in sys.argv:
        print(build_prof())
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
