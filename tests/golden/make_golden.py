#!/usr/bin/env python3
"""Generates tests/golden/* from the REAL reference compiled in oracle/_ref (run in the build container,
where /root/reference is mounted:  make -C oracle ref && python tests/golden/make_golden.py).

Outputs (all committed; nothing here is read from /root/reference at test time):
  scene_*.json            the reference's shipped scenes re-serialised (inputs), rtiow_486.json (random_scene()
                          of cmake-cpu-version under mt19937(5489)), dna_000.json (dna.py frame 0)
  kat.npz                 per-function known-answer vectors: inputs (seeded) + the reference's outputs
  aov_<scene>.npz         first-hit records at 96x54 from the reference (float shim / double cmake-cpu)
  conv_<scene>.npz        converged float images at 64x36 from the reference, two independent seeds each
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/gpu-version"
GPU_H = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_harness")
CPU_H = os.path.join(ROOT, "oracle", "_ref", "ref_cpu_harness")
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

rt = ge.load_package()
from rt_b200 import animation  # noqa: E402

AOV_W, AOV_H = 96, 54
CONV_W, CONV_H = 64, 36


def run(cmd, **kw):
    return subprocess.run(cmd, check=True, capture_output=True, text=True, **kw)


def harness_kat(kind, inp):
    with tempfile.TemporaryDirectory() as td:
        fi, fo = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        np.ascontiguousarray(inp, np.float32).tofile(fi)
        run([GPU_H, "kat_" + kind, fi, fo])
        return np.fromfile(fo, np.float32)


def with_size(scene, w, h, spp=None):
    d = json.loads(json.dumps(scene))
    # NOTE: the reference derives the camera aspect from width/height (parser.hpp:122); keep 16:9 sizes
    d["width"], d["height"] = w, h
    if spp:
        d["samples_per_pixel"] = spp
    return d


def harness_scene(mode, scene, extra=()):
    with tempfile.TemporaryDirectory() as td:
        fs, fo = os.path.join(td, "s.json"), os.path.join(td, "out.bin")
        json.dump(scene, open(fs, "w"))
        r = run([GPU_H, mode, fs] + [str(e) for e in extra[:1]] + [fo] + [str(e) for e in extra[1:]])
        return np.fromfile(fo, np.float32), r.stdout


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def make_kats(rng):
    k = {}
    n = 400
    # sphere: rays aimed near the sphere so that ~half hit; includes negative radius and origin-inside cases
    c = rng.uniform(-5, 5, (n, 3)); r = rng.uniform(0.1, 2, (n, 1)) * np.where(rng.random((n, 1)) < 0.15, -1, 1)
    o = c + unit(rng.normal(size=(n, 3))) * np.abs(r) * rng.uniform(0.2, 6, (n, 1))
    tgt = c + rng.normal(size=(n, 3)) * np.abs(r) * 0.8
    d = (tgt - o) * rng.uniform(0.3, 3, (n, 1))
    tmin = np.full((n, 1), 0.001); tmax = np.where(rng.random((n, 1)) < 0.2, rng.uniform(0.5, 3, (n, 1)), 3.4e38)
    k["sphere_in"] = np.hstack([o, d, c, r, tmin, tmax]).astype(np.float32)
    # rects
    axis = rng.integers(0, 3, (n, 1)).astype(np.float64)
    a0 = rng.uniform(-3, 0, (n, 1)); a1 = a0 + rng.uniform(0.5, 4, (n, 1)); b0 = rng.uniform(-3, 0, (n, 1)); b1 = b0 + rng.uniform(0.5, 4, (n, 1))
    kk = rng.uniform(-2, 2, (n, 1))
    o = rng.uniform(-6, 6, (n, 3))
    ta = a0 + (a1 - a0) * rng.uniform(-0.3, 1.3, (n, 1)); tb = b0 + (b1 - b0) * rng.uniform(-0.3, 1.3, (n, 1))
    tgt = np.where(axis == 0, np.hstack([ta, tb, kk]), np.where(axis == 1, np.hstack([ta, kk, tb]), np.hstack([kk, ta, tb])))
    d = (tgt - o) * rng.uniform(0.3, 3, (n, 1))
    d[::17] *= np.array([1, 0, 1])     # some axis-parallel directions (division by zero path)
    k["rect_in"] = np.hstack([axis, o, d, a0, a1, b0, b1, kk, tmin, tmax]).astype(np.float32)
    # cylinders
    rad = rng.uniform(0.2, 2, (n, 1)); zmin = rng.uniform(-2, 0, (n, 1)); zmax = zmin + rng.uniform(0.2, 3, (n, 1))
    ax = unit(rng.normal(size=(n, 3))) * rng.uniform(0.5, 2, (n, 1)); ang = rng.uniform(-180, 360, (n, 1)); tr = rng.uniform(-4, 4, (n, 3))
    o = tr + unit(rng.normal(size=(n, 3))) * rng.uniform(0.1, 8, (n, 1))
    d = (tr + rng.normal(size=(n, 3)) * 0.8 - o) * rng.uniform(0.3, 2, (n, 1))
    k["cyl_in"] = np.hstack([o, d, rad, zmin, zmax, ax, ang, tr, tmin, np.full((n, 1), 3.4e38)]).astype(np.float32)
    # scatter: type d3 p3 n3 front albedo3 fuzz ior | pad to 16 | uniforms[16]
    t = rng.integers(0, 4, (n, 1)).astype(np.float64)
    nrm = unit(rng.normal(size=(n, 3))); din = rng.normal(size=(n, 3)) * rng.uniform(0.2, 3, (n, 1))
    flip = np.sum(din * nrm, -1, keepdims=True) > 0
    nrm = np.where(flip, -nrm, nrm)                      # shading normal always faces the ray (set_face_normal)
    p = rng.uniform(-5, 5, (n, 3)); front = (rng.random((n, 1)) < 0.5).astype(np.float64)
    alb = rng.uniform(0, 1, (n, 3)); fuzz = rng.uniform(0, 1.3, (n, 1)); ior = rng.uniform(1.1, 2.5, (n, 1))
    uni = 1.0 - rng.random((n, 16)); uni[:, 12:15] = 0.1   # (0,1]; the 5th triple always lies inside the unit ball
    k["scatter_in"] = np.hstack([t, din, p, nrm, front, alb, fuzz, ior, uni]).astype(np.float32)
    # camera
    frm = rng.uniform(-10, 10, (n, 3)); at = frm + rng.normal(size=(n, 3)) * 5
    vup = np.tile([0, 1, 0], (n, 1)) + rng.normal(size=(n, 3)) * 0.1
    k["camera_in"] = np.hstack([frm, at, vup, rng.uniform(10, 90, (n, 1)), rng.uniform(0.5, 2.5, (n, 1)), rng.uniform(0, 0.3, (n, 1)),
                                rng.random((n, 2))]).astype(np.float32)
    # xform
    k["xform_in"] = np.hstack([ax, ang, tr, rng.uniform(-5, 5, (n, 3))]).astype(np.float32)
    for kind in ("sphere", "rect", "cyl", "scatter", "camera", "xform"):
        k[kind + "_out"] = harness_kat(kind, k[kind + "_in"])
    return k


def main():
    rng = np.random.default_rng(20231104)
    scenes = {}
    for name in ("sample_scene", "blue", "blue2", "basic_scene"):
        d = json.load(open(os.path.join(REF, name + ".json")))
        scenes[name] = d
        json.dump(d, open(os.path.join(HERE, name + ".json"), "w"), separators=(",", ":"))
    scenes["dna_000"] = animation.dna_frame(scenes["basic_scene"], 0)
    json.dump(scenes["dna_000"], open(os.path.join(HERE, "dna_000.json"), "w"), separators=(",", ":"))
    run([CPU_H, "dump_scene", os.path.join(HERE, "rtiow_486.json")])

    np.savez_compressed(os.path.join(HERE, "kat.npz"), **make_kats(rng))

    for name in ("sample_scene", "blue", "blue2", "dna_000"):
        a, _ = harness_scene("aov", with_size(scenes[name], AOV_W, AOV_H))
        np.savez_compressed(os.path.join(HERE, f"aov_{name}.npz"), aov=a.reshape(AOV_H, AOV_W, 24))
    with tempfile.TemporaryDirectory() as td:
        fo = os.path.join(td, "o.bin")
        run([CPU_H, "aov", str(AOV_W), str(AOV_H), fo])
        np.savez_compressed(os.path.join(HERE, "aov_rtiow_486.npz"), aov=np.fromfile(fo, np.float32).reshape(AOV_H, AOV_W, 24))

    # converged images: two independent estimates each (seed 1 / seed 2) to calibrate the tolerance
    nthreads = os.cpu_count() or 1
    for name, spp in (("sample_scene", 4096), ("blue2", 4096), ("dna_000", 1024)):
        imgs = []
        for seed in (1, 2):
            a, out = harness_scene("render", with_size(scenes[name], CONV_W, CONV_H, spp), extra=(seed, nthreads))
            imgs.append(a.reshape(CONV_H, CONV_W, 3) / spp)
            print(name, seed, out.strip())
        np.savez_compressed(os.path.join(HERE, f"conv_{name}.npz"), a=imgs[0], b=imgs[1], spp=spp)
    with tempfile.TemporaryDirectory() as td:
        imgs = []
        spp = 1024
        procs = []
        for i, skip in enumerate((0, 1000003)):
            fo = os.path.join(td, f"o{i}.bin")
            procs.append((subprocess.Popen([CPU_H, "render", str(CONV_W), str(CONV_H), str(spp), "50", str(skip), fo],
                                           stdout=subprocess.PIPE, text=True), fo))
        for p, fo in procs:
            out, _ = p.communicate()
            print("rtiow_486", out.strip())
            imgs.append(np.fromfile(fo, np.float32).reshape(CONV_H, CONV_W, 3) / spp)
        np.savez_compressed(os.path.join(HERE, "conv_rtiow_486.npz"), a=imgs[0], b=imgs[1], spp=spp)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
