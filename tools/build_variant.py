#!/usr/bin/env python3
"""Builds a tuning variant of librt_b200.so with extra -D flags into gpurun-visible build_variants/<name>.so
(select it at run time with RT_B200_LIB=build_variants/<name>.so).  Usage: build_variant.py name -DX=1 -DY=2"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
name, defs = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "build_variants"); os.makedirs(out, exist_ok=True)
cu, _ = ge._sources()
subprocess.check_call(["nvcc", "-shared"] + ge.NVCC_FLAGS + defs + ["-x", "cu"] + cu + ["-o", os.path.join(out, name + ".so"), "-lz"])
print("built", name)
