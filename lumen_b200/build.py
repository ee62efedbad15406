"""Builds lumen_b200/libb200clip.so in-tree with nvcc for sm_100a (the only target)."""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libb200clip.so"
SOURCES = ["api.cu", "preprocess.cu", "gemm_tc.cu", "vit_ops.cu", "attention_tc.cu", "topk.cu", "classify.cu", "jpeg.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--use_fast_math=false",
              "-Xcompiler", "-fPIC,-ffp-contract=off", "-Xptxas", "-v"]


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + [PKG.parent / "include" / "b200clip.h"]
    return any(p.stat().st_mtime > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    objs = []
    (PKG / "build").mkdir(exist_ok=True)
    procs = []
    for s in SOURCES:
        o = PKG / "build" / (s + ".o")
        cmd = ["nvcc", *[f for f in NVCC_FLAGS if f != "--use_fast_math=false"], "-c", str(CSRC / s), "-o", str(o)]
        procs.append((s, o, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for s, o, p in procs:
        out, _ = p.communicate()
        log.append(f"== {s}\n{out}")
        if p.returncode != 0:
            sys.stderr.write("\n".join(log))
            raise RuntimeError(f"nvcc failed on {s}")
        objs.append(str(o))
    (PKG / "build" / "ptxas.log").write_text("\n".join(log))
    subprocess.check_call(["nvcc", "-shared", "-o", str(LIB), *objs, "-lcudart"])
    if verbose:
        print("\n".join(log))
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
