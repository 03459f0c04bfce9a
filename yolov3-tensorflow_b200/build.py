"""Builds libyv3head.so (the C-ABI library of include/yv3_head.h) in-tree with nvcc for sm_100a.

    python -m yolov3_tensorflow_b200.build        # or __graft_entry__.build()

Every translation unit is compiled with -fmad=false: the reference rounds each operator once
(TensorFlow/Eigen elementwise kernels), and decisions (argmax, >=, < iou_thresh, IoU > nms_thresh)
must see those roundings.  The kernels are memory-bound; the lost FMA fusion costs nothing.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libyv3head.so")
SOURCES = ["yv3_api.cu", "assign.cu", "loss_dense.cu", "decode.cu", "postprocess.cu", "fused_sweep.cu", "peer_reduce.cu", "map_eval.cu",
           "static_scratch.cu"]
HEADERS = ["yv3_common.cuh", "yv3_kernels.cuh", "tile_ops.cuh", "sweep_parts.cuh", os.path.join("..", "..", "include", "yv3_head.h")]

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
] + os.environ.get("YV3_EXTRA_NVCC", "").split()        # developer experiments (-D switches); part of the fingerprint


# postprocess.cu: the sort kernel tail-launches the queue-driven NMS kernels (CUDA dynamic parallelism)
RDC = {"postprocess.cu": ["-rdc=true"]}


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libyv3head.so cannot be built")


def _fingerprint():
    """Content hash of every source, header and flag: mtimes do not survive the copy to the GPU box."""
    import hashlib
    h = hashlib.sha256()
    h.update((" ".join(NVCC_FLAGS) + repr(sorted(RDC.items()))).encode())
    for d in [os.path.join(CSRC, s) for s in SOURCES + HEADERS]:
        with open(d, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


SWEEP_SOURCES = ["fused_sweep.cu", "loss_dense.cu", "decode.cu", "sweep_parts.cuh", "tile_ops.cuh", "yv3_common.cuh",
                 "yv3_kernels.cuh"]


def sweep_sources_hash():
    """Content hash of the files the HBM-bound sweep kernels are made of: profiles/traffic.json carries it, and bench.py
    marks its DRAM-traffic figures stale when these files changed since the capture (a commit elsewhere does not)."""
    import hashlib
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for name in SWEEP_SOURCES:
        with open(os.path.join(CSRC, name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def _stale():
    stamp = os.path.join(LIB_DIR, "build.hash")
    if not os.path.exists(LIB_PATH) or not os.path.exists(stamp):
        return True
    with open(stamp) as f:
        return f.read().strip() != _fingerprint()


class NvccMissing(RuntimeError):
    pass


def build_library(force=False, verbose=False):
    """Compile the library if it is missing or its sources changed.  Returns its path.  Safe when several processes
    (the ranks of one torchrun) call it at once: the build runs under an exclusive file lock, objects and the library
    are written under temporary names and moved into place, and a rank that waited for the lock finds a fresh stamp."""
    import fcntl
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    with open(os.path.join(LIB_DIR, ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():          # another process built it while this one waited
                return LIB_PATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose):
    try:
        nvcc = _nvcc()
    except RuntimeError as e:
        raise NvccMissing(str(e))
    objs = []
    log = []
    tag = ".tmp%d" % os.getpid()
    for src in SOURCES:
        obj = os.path.join(LIB_DIR, src.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + RDC.get(src, []) + ["-c", os.path.join(CSRC, src), "-o", obj + tag]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            sys.stderr.write(log[-1])
            raise RuntimeError("nvcc failed on %s" % src)
        os.replace(obj + tag, obj)
        objs.append(obj)
    # nvcc device-links the relocatable object (postprocess.cu launches kernels from the device) against cudadevrt
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH + tag] + objs + ["-lcudadevrt", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        sys.stderr.write(log[-1])
        raise RuntimeError("link failed")
    os.replace(LIB_PATH + tag, LIB_PATH)
    with open(os.path.join(LIB_DIR, "build.log"), "w") as f:
        f.write("\n".join(log))
    with open(os.path.join(LIB_DIR, "build.hash" + tag), "w") as f:
        f.write(_fingerprint())
    os.replace(os.path.join(LIB_DIR, "build.hash" + tag), os.path.join(LIB_DIR, "build.hash"))
    if verbose:
        print("\n".join(log))
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
