#!/usr/bin/env python
"""Build a kernel variant of libgs2d.so for A/B runs on the GPU box:
    python profiles/build_variant.py NAME -DGS2D_BWD_ROWS=4 [-DGS2D_TILE_MIN_BLOCKS=12 ...]
 -> gsvc_b200/lib/variants/libgs2d_NAME.so (git-ignored, travels with gpurun); run with
    GS2D_LIB=gsvc_b200/lib/variants/libgs2d_NAME.so python bench.py ..."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gsvc_b200 import build as B  # noqa: E402

name, defs = sys.argv[1], sys.argv[2:]
out_dir = os.path.join(B.LIBDIR, "variants")
os.makedirs(out_dir, exist_ok=True)
objs = []
for src in B.SOURCES:
    obj = os.path.join(out_dir, f"{name}_{src.replace('.cu', '.o')}")
    subprocess.run([B._nvcc(), *B.NVCC_FLAGS, *defs, "-c", os.path.join(B.CSRC, src), "-o", obj], check=True)
    objs.append(obj)
lib = os.path.join(out_dir, f"libgs2d_{name}.so")
subprocess.run([B._nvcc(), "-shared", "-o", lib, *objs], check=True)
for o in objs:
    os.remove(o)
print(lib)
