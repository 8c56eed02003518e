"""In-tree build of libtaco_b200.so (hand-written sm_100a kernels + C ABI) with nvcc.

``python -m tacotron_prosody_multi_gpu_b200.build`` or ``build()``; nvcc cross-compiles for
sm_100a without a GPU.  The .so is git-ignored but travels with the gpurun snapshot.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "libtaco_b200.so")
SOURCES = ["api.cu", "decoder.cu", "decoder3.cu", "frame_proj.cu", "conv.cu", "conv_tc.cu", "encoder.cu", "refenc.cu",
           "train_gemm.cu", "train_ops.cu", "train_attn.cu", "train_attn_fused.cu", "train_misc.cu", "train_loops.cu", "griffin_lim.cu"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the CUDA extension cannot be built (there is no CPU fallback)")
    return exe


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inl"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "taco_b200.h"))
    headers.append(os.path.join(os.path.dirname(HERE), "include", "taco_b200_train.h"))
    headers.append(os.path.join(os.path.dirname(HERE), "include", "taco_b200_audio.h"))
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for s in sources:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r

    if jobs:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, r in ex.map(compile_one, jobs):
                if verbose or r.returncode != 0:
                    sys.stderr.write("== %s\n%s%s\n" % (os.path.basename(src), r.stdout, r.stderr))
                if r.returncode != 0:
                    raise RuntimeError("nvcc failed on %s" % src)
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in sources]
    if force or jobs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-lcudart", "-lcuda"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
