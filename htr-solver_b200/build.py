"""Builds libhtr_b200_<Mix>.so (one per mixture) in-tree with nvcc for sm_100a.

    python htr-solver_b200/build.py [--mix AirMix ...] [--force]
"""
import argparse, os, subprocess, sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
MIXES = ["ConstPropMix", "AirMix"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr",
         "-ccbin", "/usr/bin/g++"]


def lib_path(mix):
    return os.path.join(HERE, f"libhtr_b200_{mix}.so")


def sources():
    out = []
    for root, _, files in os.walk(CSRC):
        out += [os.path.join(root, f) for f in files if f.endswith((".cu", ".cuh", ".h", ".inc"))]
    out.append(os.path.join(os.path.dirname(HERE), "include", "htr_b200.h"))
    return out


def needs_build(mix):
    lib = lib_path(mix)
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(s) > t for s in sources())


def build(mixes=MIXES, force=False, extra=(), verbose=False):
    for mix in mixes:
        if not force and not needs_build(mix):
            continue
        cmd = [NVCC] + ARCH + FLAGS + list(extra) + [f"-DHTR_MIX_{mix}", os.path.join(CSRC, "htr_kernels.cu"),
                                                     "-o", lib_path(mix)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return [lib_path(m) for m in mixes]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mix", nargs="*", default=MIXES)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-v", action="store_true")
    ap.add_argument("extra", nargs="*")
    a = ap.parse_args()
    print(build(a.mix, a.force, a.extra, a.v))
