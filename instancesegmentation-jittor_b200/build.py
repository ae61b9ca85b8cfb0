"""Build libroipath_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python instancesegmentation-jittor_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels to the GPU box with the
gpurun snapshot.  Flags:
  -gencode arch=compute_100a,code=sm_100a   B200 only, no PTX fallback for other archs
  --fmad=false                              no implicit FMA contraction: the bit-exact paths
                                            (NMS IoU, mask sigmoid/bilinear) must round exactly
                                            like the gcc -ffp-contract=off oracle; fmaf() is
                                            written explicitly where wanted
  -lineinfo                                 ncu source page maps to these files
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libroipath_b200.so")
SOURCES = ["roi_align.cu", "roi_fwd_tc.cu", "roi_nhwc.cu", "nms.cu", "yolact.cu", "masks.cu", "mask_loss.cu", "rpn.cu", "box_post.cu", "roi_misc.cu", "capi.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo", "--fmad=false",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O2",
    "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
]


def _deps():
    out = []
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            if f.endswith((".cu", ".cuh", ".h")):
                out.append(os.path.join(root, f))
    return out


def build(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    if not force and os.path.exists(LIB):
        t = os.path.getmtime(LIB)
        if all(os.path.getmtime(d) <= t for d in _deps()):
            return LIB
    if not os.path.exists(NVCC):
        raise RuntimeError("nvcc not found at %s; cannot build libroipath_b200.so" % NVCC)
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    procs = []
    for s in srcs:
        o = os.path.join(HERE, "build", os.path.basename(s) + ".o")
        objs.append(o)
        cmd = [NVCC] + FLAGS + os.environ.get("RP_EXTRA_NVCC", "").split() + ["-c", s, "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for s, p in procs:
        out, _ = p.communicate()
        log.append(out)
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed on %s" % s)
    with open(os.path.join(HERE, "build", "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                  "-Xcompiler", "-fPIC", "-lcudart_static"]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
