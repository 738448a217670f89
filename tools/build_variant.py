#!/usr/bin/env python3
"""Builds a tuning variant of librt_b200.so with extra -D flags into gpurun-visible build_variants/<name>.so
(select it at run time with RT_B200_LIB=build_variants/<name>.so).  Usage: build_variant.py name [rev=<git rev>] -DX=1 -DY=2
rev=<git rev>: build the kernels as they were at that commit (A/B on the same GPU box)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
name, defs = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "build_variants"); os.makedirs(out, exist_ok=True)
cu, _ = ge._sources()
if defs and defs[0].startswith("rev="):
    rev, defs = defs[0][4:], defs[1:]
    tree = os.path.join(out, "src_" + name); os.makedirs(tree, exist_ok=True)
    subprocess.check_call(["git", "--work-tree=" + tree, "checkout", rev, "--", "ray-tracing-in-cuda_b200/csrc", "include"], cwd=ROOT)
    subprocess.check_call(["git", "reset", "-q"], cwd=ROOT)       # the checkout above touched the index only for those paths
    cu = [os.path.join(tree, os.path.relpath(f, ROOT)) for f in cu]
    defs = ["-I" + os.path.join(tree, "include")] + defs
subprocess.check_call(["nvcc", "-shared"] + ge.NVCC_FLAGS + defs + ["-x", "cu"] + cu + ["-o", os.path.join(out, name + ".so"), "-lz"])
print("built", name)
