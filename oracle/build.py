"""Build recipe for the oracle's C restatement (TEST INFRASTRUCTURE).

Compiles oracle/host_kernels.c into oracle/_build/liboracle.so with gcc.
The reference itself is Rust (moose/Cargo.toml) and there is no cargo/rustc
in this image, so there is no oracle/_ref to build: the oracle is a port.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_kernels.c")
OUT_DIR = os.path.join(HERE, "_build")
OUT = os.path.join(OUT_DIR, "liboracle.so")


def build(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    if (not force and os.path.exists(OUT)
            and os.path.getmtime(OUT) >= os.path.getmtime(SRC)):
        return OUT
    cmd = ["gcc", "-O3", "-fPIC", "-shared", "-fopenmp", "-std=gnu11",
           "-Wall", "-Wextra", "-o", OUT, SRC]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
