"""Builds libloam_b200.so (sm_100a only) and the host-only synthetic sweep generator, in-tree."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libloam_b200.so")


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_cuda(force=False, verbose=False):
    csrc = os.path.join(HERE, "csrc")
    srcs = [os.path.join(csrc, f) for f in os.listdir(csrc)] + [
        os.path.join(ROOT, "include", "loam_b200.h"), os.path.join(ROOT, "include", "loam_b200_debug.h"), os.path.join(ROOT, "include", "loam", "detmath.h")]
    if not force and not _newer(LIB, srcs):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
           "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
           "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-shared",
           "-I", os.path.join(ROOT, "include"), "-I", csrc,
           os.path.join(csrc, "loam_b200.cu"), "-o", LIB, "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
    for var in ("LOAM_LO_MAXREG", "LOAM_LM_MAXREG", "LOAM_RF_SLICES", "LOAM_RF_CLUSTERS", "LOAM_LM_THREADS", "LOAM_LM_KW", "LOAM_LO_CELL_CORNER", "LOAM_LO_CELL_SURF"):   # dev builds: register caps of the two persistent kernels
        if os.environ.get(var):
            cmd.insert(1, "-D%s=%s" % (var, os.environ[var]))
    if os.environ.get("LOAM_LIB_OUT"):
        cmd[cmd.index("-o") + 1] = os.environ["LOAM_LIB_OUT"]
    if os.environ.get("LOAM_PHASE_CLOCKS"):   # dev builds only: clock64 stamps inside the persistent LM kernels
        cmd.insert(1, "-DLOAM_PHASE_CLOCKS")
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


def build_all(force=False):
    sys.path.insert(0, ROOT)
    import importlib
    importlib.import_module("loam_velodyne-1_b200.synth").build()
    return build_cuda(force=force)


if __name__ == "__main__":
    print(build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv))
