"""Build the two native libraries in-tree (they travel to the GPU box with the snapshot).

libceres_b200_schur.so  CUDA kernels + C ABI (include/ceres_b200_schur.h), sm_100a only
libsynthetic_bal.so     host-only synthetic BAL problems (measurement / test support)
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_CUDA = os.path.join(HERE, "libceres_b200_schur.so")
LIB_BAL = os.path.join(HERE, "libsynthetic_bal.so")
LIB_LM = os.path.join(HERE, "libceres_b200_lm.so")

_CUDA_SRC = ["c_api.cu", "c_api_solve.inc", "c_api_entry.inc", "c_api_lm.inc", "lm_kernels.cuh", "snavely_model.h", "kernels.cuh", "camera_kernels.cuh",
             "cg_kernels.cuh", "cg_step.cuh", "stream_kernels.cuh", "solver_state.h", "flatten.h",
             "c_api_explicit.inc", "explicit_kernels.cuh", "explicit_plan.h", "multi_gpu.h"]


def _stale(out, srcs):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in srcs)


def nvcc_path():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build_cuda(force=False, verbose=False):
    srcs = [os.path.join(CSRC, f) for f in _CUDA_SRC]
    srcs.append(os.path.join(HERE, "..", "include", "ceres_b200_schur.h"))
    if not force and not _stale(LIB_CUDA, srcs):
        return LIB_CUDA
    cmd = [nvcc_path(), "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a",
           "-lineinfo", "-Xcompiler", "-fPIC", "-shared", "-o", LIB_CUDA,
           os.path.join(CSRC, "c_api.cu"), "-ldl"]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB_CUDA


def build_bal(force=False):
    src = os.path.join(CSRC, "synthetic_bal.cc")
    if not force and not _stale(LIB_BAL, [src, os.path.join(CSRC, "snavely_model.h"), os.path.join(CSRC, "bal_io.h")]):
        return LIB_BAL
    # x86-64-v3 (not -march=native): the .so is built here and run on the GPU box
    subprocess.run(["g++", "-O3", "-march=x86-64-v3", "-std=c++17", "-fopenmp", "-fPIC", "-shared",
                    "-o", LIB_BAL, src], check=True, cwd=CSRC)
    return LIB_BAL


CLI = os.path.join(HERE, "bin", "bundle_adjuster")


def build_cli(force=False):
    """tools/bundle_adjuster.cc: the reference's examples/bundle_adjuster flow over the C ABI."""
    src = os.path.join(HERE, "..", "tools", "bundle_adjuster.cc")
    deps = [src, os.path.join(CSRC, "bal_io.h"), os.path.join(CSRC, "snavely_model.h"),
            os.path.join(HERE, "..", "include", "ceres_b200_schur.h"), LIB_CUDA]
    build_cuda()
    if not force and not _stale(CLI, deps):
        return CLI
    os.makedirs(os.path.dirname(CLI), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", CLI, src, "-L" + HERE, "-lceres_b200_schur",
                    "-Wl,-rpath,$ORIGIN/..", "-ldl", "-lpthread"], check=True)
    return CLI


def build_all(force=False):
    return build_cuda(force), build_bal(force), build_cli(force)


if __name__ == "__main__":
    print(build_all(force=True))
