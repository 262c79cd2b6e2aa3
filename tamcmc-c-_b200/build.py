"""In-tree build of libtamcmc_gpu.so (sm_100a only; nvcc cross-compiles without a GPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libtamcmc_gpu.so")
SOURCES = ["tamcmc_gpu.cu", "tg_outputs.cpp", "tg_frontend.cpp"]
HEADERS = ["tg_internal.cuh", "tg_prologue.cuh", "tg_eval.cuh", "tg_farfield.cuh", "tg_adapt.cuh", "tg_sampler.cuh",
           os.path.join("..", "..", "include", "tamcmc_gpu.h"), os.path.join("..", "..", "include", "tamcmc_outputs.h"),
           os.path.join("..", "..", "include", "tamcmc_frontend.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    for f in SOURCES + HEADERS:
        p = os.path.join(CSRC, f)
        if os.path.exists(p) and os.path.getmtime(p) > t:
            return True
    return False


def build_library(force=False, verbose=False):
    """Compile csrc/*.cu -> csrc/libtamcmc_gpu.so.  Returns the library path."""
    if not force and not needs_build():
        return LIB_PATH
    extra = os.environ.get("TG_NVCC_EXTRA", "").split()            # developer builds (e.g. -DTG_EVAL_MINBLOCKS=3)
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES]
    env = dict(os.environ)
    # the image exports CC/CXX wrappers; nvcc should use the system host compiler
    if os.path.exists("/usr/bin/g++"):
        cmd[1:1] = ["-ccbin", "/usr/bin/g++"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout)
    if verbose:
        print(res.stdout)
    return LIB_PATH
