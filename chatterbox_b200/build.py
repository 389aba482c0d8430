"""In-tree build of the C-ABI library `chatterbox_b200/libcbx.so` (sm_100a only).

`nvcc` cross-compiles without a GPU, so this runs in the build container; the built `.so` travels
to the GPU box with the repo snapshot (git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcbx.so")
SOURCES = ["cbx_core.cu", "t3_decode.cu", "t3_prefill.cu", "t3_batch.cu", "t3_front.cu", "s3_gemm.cu", "s3_gemm_ws.cu", "s3_ffn.cu", "s3_attn.cu", "s3_attn_tc.cu", "s3_cfm.cu", "s3_front.cu", "s3_hift.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = _sources() + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "cbx.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    flags = list(NVCC_FLAGS)
    if verbose:
        flags += ["-Xptxas", "-v"]
    flags += os.environ.get("CBX_NVCC_EXTRA", "").split()
    cmd = [nvcc] + flags + ["-o", LIB] + _sources()
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose:
        sys.stderr.write(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
