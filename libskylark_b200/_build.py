"""In-tree build of libcskylark_b200.so (nvcc, sm_100a only; no torch in the link)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.environ.get("SKB_SO") or os.path.join(HERE, "libcskylark_b200.so")   # SKB_SO: A/B builds
SOURCES = ["capi.cu", "hash.cu", "dense.cu", "fjlt.cu"]
HEADERS = ["rng.cuh", "cr_math.cuh", "cr_tables.cuh", "skb_internal.h", os.path.join("..", "..", "include", "skylark_b200.h")]


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def is_stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into one shared library next to this file."""
    if not force and not is_stale():
        return SO
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libcskylark_b200.so")
    cmd = [nvcc, "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
           "-Xcompiler", "-fPIC", "-shared", "-o", SO] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    if os.environ.get("SKB_NVCC_FLAGS"):        # A/B builds (with SKB_SO), e.g. -DSKB_LIBM_FAST
        cmd[1:1] = os.environ["SKB_NVCC_FLAGS"].split()
    subprocess.check_call(cmd)
    return SO


if __name__ == "__main__":
    print(build(force=True, verbose=True))
