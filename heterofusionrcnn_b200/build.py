"""Build recipe for libhfops.so (hand-written sm_100a CUDA behind the C-ABI of include/hfops.h).

    python -m heterofusionrcnn_b200.build [--force]

nvcc cross-compiles for sm_100a without a GPU.  The library is built IN-TREE
(heterofusionrcnn_b200/libhfops.so) so that it travels with the source tree.
"""
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libhfops.so")
SOURCES = ["api.cu", "sampling.cu", "fps_bucket.cu", "fps_multi.cu", "grouping.cu", "spatial.cu", "interpolate.cu", "cropping.cu", "bev_iou.cu", "backward.cu", "topk.cu", "glue.cu", "prob_sample.cu", "lidar_input.cu", "pc_sample.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; libhfops.so cannot be built")
    return exe


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(PKG_DIR, "..", "include", "hfops.h"))
    return any(os.path.getmtime(d) > built for d in deps if os.path.exists(d))


def source_id():
    """12 hex digits over every file the library is built from: hf_version() carries it, so a profile or a
    measurement can name the exact build it was taken from (bench.py refuses ncu traffic of another build)."""
    import hashlib

    h = hashlib.sha256()
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(PKG_DIR, "..", "include", "hfops.h")]
    for f in files:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:12]


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link libhfops.so.  Returns the path."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = _nvcc()
    NVCC_FLAGS = globals()["NVCC_FLAGS"] + ['-DHF_BUILD_ID="%s"' % source_id()]
    objdir = os.path.join(PKG_DIR, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, out))
        if verbose:
            sys.stderr.write(out)
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH] + objs
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("link of libhfops.so failed:\n%s" % res.stdout)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
