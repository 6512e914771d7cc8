"""Build libb200tsdf.so in-tree with nvcc for sm_100a (no torch involved).

`python -m mast3r_slam_quality_dualtsdf_b200.build [--force]`
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200tsdf.so")
SOURCES = ["volume.cu", "integrate.cu", "sweep.cu", "mc.cu", "weld.cu", "local.cu", "quality.cu"]
HEADERS = ["common.cuh", "ray.cuh", "march_fast.cuh", "fuse.cuh", "fuse_fast.cuh", "march_key.h",
           "sim3.cuh", "mc_tables.h", os.path.join("..", "..", "include", "b200tsdf.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    "-Xptxas", "-v",
]


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: libb200tsdf cannot be built (no CPU fallback exists)")
    return p


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    cmd = [nvcc_path()] + NVCC_FLAGS + ["-o", LIB] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libb200tsdf.so")
    log = os.path.join(HERE, "csrc", "ptxas.log")
    with open(log, "w") as f:
        f.write(r.stderr)
    return LIB


def build_variant(name, defines):
    """A/B build of the same library with extra -D flags -> build/variants/lib_<name>.so
    (select it with the B200TSDF_LIB environment variable; scripts/ab_variants.sh)."""
    out_dir = os.path.join(os.path.dirname(HERE), "build", "variants")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, f"lib_{name}.so")
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    cmd = [nvcc_path()] + [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")] + \
        [f"-D{d}" for d in defines] + ["-o", out] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"nvcc failed building variant {name}")
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
