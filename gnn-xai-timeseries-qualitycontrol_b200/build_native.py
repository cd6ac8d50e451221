"""In-tree build of libgnnqc.so (nvcc, sm_100a only).

`python build_native.py [-v] [-f]` or `build_native.build()`; the shared object lands in
gnnqc/ (in-tree) so that it travels with the repository snapshot.  This file must not import
the gnnqc package (which refuses to import without the library).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "gnnqc", "libgnnqc.so")
OBJ_DIR = os.path.join(CSRC, "build")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-Wall", "--expt-relaxed-constexpr"]

SOURCES = ["tf_bundle.cc", "model_spec.cc", "device_model.cu", "capi.cu", "fma_probe.cu", "no_alloc.cu", "ig.cu", "soilnet.cu", "train.cu", "series.cu", "tfrecord.cc", "umma_selftest.cu", "lstm_tc.cu", "lstm_tc_bwd.cu", "host_scoring.cu", "metrics.cu", "soil_series.cu"]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def build(verbose=False, force=False, variant=None, defines=()):
    """variant / defines: an A/B build (e.g. variant="p3", defines=["-DGNNQC_POLY_EXPS=3"]) lands in
    gnnqc/libgnnqc_<variant>.so with its own object directory; GNNQC_LIB selects it at import."""
    global OBJ_DIR, OUT
    if variant:
        OBJ_DIR = os.path.join(CSRC, "build", "variant_" + variant)
        OUT = os.path.join(HERE, "gnnqc", "libgnnqc_%s.so" % variant)
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.normpath(os.path.join(HERE, "..", "include", "gnnqc.h")))
    objs = []
    for name in SOURCES:
        src = os.path.join(CSRC, name)
        if not os.path.exists(src):
            continue
        obj = os.path.join(OBJ_DIR, os.path.splitext(name)[0] + ".o")
        objs.append(obj)
        if force or _newer([src] + headers, obj):
            cmd = [NVCC] + ARCH + COMMON + list(defines) + ["-Xptxas", "-v", "-c", src, "-o", obj]
            res = subprocess.run(cmd, capture_output=True, text=True)
            with open(obj + ".log", "w") as f:
                f.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
            if verbose or res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
            if res.returncode != 0:
                raise RuntimeError("nvcc failed on %s" % name)
    if force or _newer(objs, OUT):
        cmd = [NVCC] + ARCH + ["-shared", "-cudart", "static", "-o", OUT] + objs
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    var = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--variant=")), None)
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv, variant=var, defines=[a for a in sys.argv if a.startswith("-D")]))
