"""Builds cpu-ray-tracing-implementation_b200/libb200rt.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libb200rt.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",  # reference-order binary64: no FMA contraction (DESIGN.md "Numerics")
    "-Xcompiler", "-fPIC,-ffp-contract=off",
    "-shared",
]


def sources():
    return [os.path.join(CSRC, "api.cu")]


def deps():
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    out.append(os.path.join(HERE, "..", "include", "b200rt.h"))
    return out


def build(force=False, verbose=False, out=None):
    global OUT
    if out:
        OUT, force = out, True
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps()):
        return OUT
    nvcc = os.environ.get("NVCC", "nvcc")
    extra = os.environ.get("B200RT_NVCC_EXTRA", "").split()  # experiments only (e.g. -DB200_SHADE_MINB=6)
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + sources() + ["-o", OUT]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libb200rt.so")
    if verbose:
        print(r.stderr)
    return OUT


if __name__ == "__main__":
    alt = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--out=")]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, out=alt[0] if alt else None))
