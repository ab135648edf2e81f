"""Builds libmcov_b200.so in-tree with nvcc for sm_100a (no torch headers, no JIT cache).

The .so is git-ignored but travels to the GPU box with the gpurun snapshot."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmcov_b200.so")
SOURCES = ["mcov_api.cu", "mcov_dec.cu", "mcov_dec_cta.cu", "mcov_super.cu", "mcov_extras.cu", "mcov_replay.cu", "mcov_aux.cu"]
HEADERS = ["mcov_common.cuh", "mcov_dec_shared.cuh", "mcov_dt.cuh", "mcov_internal.h", os.path.join("..", "..", "include", "mcov.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libmcov_b200.so cannot be built")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compiles every .cu to an object in parallel (one nvcc per file), then links the shared library."""
    if not force and not needs_build():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    extra = os.environ.get("MCOV_NVCC_EXTRA", "").split()      # tuning experiments (-DMCOV_SUPER_MINB=2 ...)
    nvcc = _nvcc()
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + extra + (["-Xptxas", "-v"] if verbose else [])
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(objdir, src[:-3] + ".o")
        return obj, subprocess.run([nvcc] + flags + ["-c", os.path.join(CSRC, src), "-o", obj],
                                   capture_output=True, text=True)

    with ThreadPoolExecutor(len(SOURCES)) as pool:
        results = list(pool.map(compile_one, SOURCES))
    for obj, res in results:
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("nvcc failed building libmcov_b200.so")
        if verbose:
            sys.stderr.write(res.stderr)
    tmp = LIB + ".tmp.%d" % os.getpid()                        # link aside, then rename: the library is never half-written
    res = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC"] +
                         [obj for obj, _ in results] + ["-o", tmp], capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libmcov_b200.so")
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose="-v" in sys.argv))
