"""Build recipe for libmoose_b200.so (sm_100a only, in-tree).

    python -m moose_b200.build [--force]

Each source is compiled with
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17
and linked into moose_b200/libmoose_b200.so (cudart linked statically, libcuda
dynamically for the TMA descriptor encoder).  nvcc cross-compiles without a GPU.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libmoose_b200.so")
SOURCES = ["core.cpp", "hostcrypto.cpp", "api.cpp", "rep_api.cpp", "rep_math.cpp", "rep_shape.cpp", "executor.cpp", "party.cpp", "nccl_transport.cpp", "peer_transport.cu", "party_fused.cu", "elementwise.cu", "sample.cu",
           "rep.cu", "bits.cu", "bits_fused.cu", "math_fused.cu", "host_misc.cu", "shape_ops.cu", "storage.cu", "gemm_imad.cu", "gemm_limb.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function", "-x", "cu"] + os.environ.get("MOOSE_B200_NVCC_FLAGS", "").split()


def _deps_mtime() -> float:
    m = 0.0
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for f in os.listdir(root):
            if f.endswith((".h", ".cuh", ".inc")):
                m = max(m, os.path.getmtime(os.path.join(root, f)))
    return m


def _compile(src: str, force: bool, hdr_mtime: float, verbose: bool) -> str:
    s = os.path.join(CSRC, src)
    o = os.path.join(OBJ, src + ".o")
    if (not force and os.path.exists(o) and os.path.getmtime(o) >= os.path.getmtime(s)
            and os.path.getmtime(o) >= hdr_mtime):
        return o
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    return o


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hdr = _deps_mtime()
    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(lambda s: _compile(s, force, hdr, verbose), SOURCES))
    if (force or not os.path.exists(LIB)
            or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs)):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-lcuda", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
