#!/usr/bin/env python
"""Build a variant of libceres_b200_schur.so with extra -D flags for A/B kernel experiments.

    python tools/build_variant.py NAME -DCB200_SOMETHING=1 ...   ->  build/variants/NAME.so
    CB200_LIBRARY=build/variants/NAME.so python tools/profile_op.py ...
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "ceres-solver-ceres-solver_b200", "csrc")


def main():
    name, flags = sys.argv[1], sys.argv[2:]
    out_dir = os.path.join(ROOT, "build", "variants")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, name + ".so")
    cmd = ["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xptxas", "-v",
           "-Xcompiler", "-fPIC", "-shared", "-o", out, os.path.join(CSRC, "c_api.cu"), "-ldl"] + flags
    r = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    lines = (r.stdout + r.stderr).splitlines()
    want = os.environ.get("SHOW", "point_stream_kernel")
    for i, l in enumerate(lines):
        if "error" in l or ("Compiling entry" in l and want in l):
            print("\n".join(lines[i:i + 4]))
    print(out if r.returncode == 0 else "BUILD FAILED")
    sys.exit(r.returncode)


if __name__ == "__main__":
    main()
