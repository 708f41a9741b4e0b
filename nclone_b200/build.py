"""Build libnppsim.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libnppsim.so")
SOURCES = ["nppsim_kernels.cu", "map_compile_device.cu", "map_compiler.cpp"]
HEADERS = ["geom.cuh", "pool.cuh", "entities.cuh", "ninja.cuh", "sim_variant.inc", "pow2.cuh", "pow2_exact.cuh", "pow2_tables.h", "map_blob.h", "tile_tables.h", "../../include/nppsim.h"]

# fp64 discipline: no FMA contraction, IEEE sqrt/div (nvcc's defaults for
# double; stated explicitly), no fast-math.
NVCC_FLAGS = [
    "-shared", "-Xcompiler", "-fPIC", "-std=c++17", "-O3",
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "--fmad=false", "--prec-div=true", "--prec-sqrt=true",
    "-Xcompiler", "-ffp-contract=off",
]


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found; libnppsim.so cannot be built")
    return p


def have_nvcc():
    return bool(shutil.which("nvcc")) or os.path.exists("/usr/local/cuda/bin/nvcc")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    for f in SOURCES + HEADERS + ["../build.py"]:
        p = os.path.join(CSRC, f)
        if os.path.exists(p) and os.path.getmtime(p) > t:
            return True
    return False


def build(force=False, verbose=False, extra=()):
    if not force and not needs_build():
        return LIB
    cmd = [nvcc_path()] + NVCC_FLAGS + list(extra) + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas"); cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
