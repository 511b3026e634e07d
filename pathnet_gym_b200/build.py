"""Build libpathnet_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libpathnet_b200.so")
SOURCES = ["pn_api.cu", "pn_net_kernels.cu", "pn_rmsprop.cu", "pn_genotype.cu", "pn_noalloc.cu", "pn_conv0_tc.cu", "pn_lin_tc.cu", "pn_conv_mid.cu", "pn_conv_mid_tc.cu", "pn_rollout.cu", "pn_peer.cu", "pn_lstm.cu", "pn_gemm_tc.cu"]
HEADERS = [os.path.join(CSRC, "pn_internal.cuh"), os.path.join(CSRC, "pn_tc.cuh"), os.path.join(CSRC, "pn_rms_core.cuh"), os.path.join(ROOT, "include", "pathnet_b200.h")]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    out = [os.path.join(CSRC, s) for s in SOURCES]
    missing = [p for p in out if not os.path.exists(p)]
    if missing:
        raise RuntimeError("listed CUDA sources are missing: %s" % ", ".join(missing))
    return out


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in sources() + HEADERS + [os.path.abspath(__file__)])


def build_library(force=False, verbose=False):
    if not force and not is_stale():
        return LIB
    cmd = [_nvcc(), "-shared", "-o", LIB, "-std=c++17", "-O3", "-lineinfo",
           "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC,-fvisibility=hidden", "-cudart", "static",
           "-I", os.path.join(ROOT, "include"), "-I", CSRC]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    # PN_BUILD_DEFINES="PN_TC_DEBUG_BUILD": compile the role-timing instrumentation of the tensor-core kernels in
    # (read with PN_TC_DEBUG=<mode> at run time; off in the product build: predicated-off instructions still take issue slots)
    cmd += ["-D" + d for d in os.environ.get("PN_BUILD_DEFINES", "").split() if d]
    cmd += sources()
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed (exit %d)" % r.returncode)
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose="-v" in sys.argv))
