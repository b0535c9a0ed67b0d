"""Builds libhtr_b200_<Mix>.so (one per mixture) in-tree with nvcc for sm_100a.

    python htr-solver_b200/build.py [--mix AirMix ...] [--force]
"""
import argparse, os, subprocess, sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
MIXES = ["ConstPropMix", "AirMix", "CH41StMix"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr",
         "-ccbin", "/usr/bin/g++"]


# variants: "" = production (FMA contraction on, nvcc default, like the reference's own GPU build);
# "strict" = -fmad=false -DHTR_EXACT_POW: no contraction, libdevice pow -- operation-for-operation
# comparable with the reference's CPU tasks; used by the parity tests to separate rounding
# from logic.
VARIANTS = {"": [], "strict": ["-fmad=false", "-DHTR_EXACT_POW", "-DHTR_FMAD_OFF"]}
# tuning experiments: HTR_EXP="name1:-DA=1,-DB=2;name2:-DC=3" adds variants built on request (--variants name1 name2)
for _e in filter(None, os.environ.get("HTR_EXP", "").split(";")):
    _n, _f = _e.split(":", 1)
    VARIANTS[_n] = [x for x in _f.split(",") if x]


def lib_path(mix, variant=""):
    return os.path.join(HERE, f"libhtr_b200_{mix}{'_' + variant if variant else ''}.so")


def sources():
    out = []
    for root, _, files in os.walk(CSRC):
        out += [os.path.join(root, f) for f in files if f.endswith((".cu", ".cuh", ".h", ".inc"))]
    out.append(os.path.join(os.path.dirname(HERE), "include", "htr_b200.h"))
    return out


def needs_build(mix, variant=""):
    lib = lib_path(mix, variant)
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(s) > t for s in sources())


def build(mixes=MIXES, force=False, extra=(), verbose=False, variants=("", "strict")):
    procs = []
    for mix in mixes:
        for var in variants:
            if not force and not needs_build(mix, var):
                continue
            cmd = [NVCC] + ARCH + FLAGS + VARIANTS[var] + list(extra) + [f"-DHTR_MIX_{mix}", os.path.join(CSRC, "htr_kernels.cu"),
                                                                          "-o", lib_path(mix, var)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
                print(" ".join(cmd))
            procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, cmd)
    return [lib_path(m, v) for m in mixes for v in variants]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mix", nargs="*", default=MIXES)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-v", action="store_true")
    ap.add_argument("--variants", nargs="*", default=["", "strict"])
    ap.add_argument("extra", nargs="*")
    a = ap.parse_args()
    print(build(a.mix, a.force, a.extra, a.v, a.variants))
