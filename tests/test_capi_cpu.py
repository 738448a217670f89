"""CPU tests of the drop-in boundary: librt_b200.so loads without a GPU, exports every symbol include/rt.h declares,
fails loudly (no CPU fallback) when there is no device, and its host-side image writers/reader are correct."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import __graft_entry__ as ge
from rt_b200 import Renderer, RtError, animation, capi, renderer

ROOT = ge.ROOT


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(capi.LIB_PATH):
        ge.build()
    return capi.load_library()


def test_header_and_library_export_the_same_symbols(lib):
    hdr = open(os.path.join(ROOT, "include", "rt.h")).read()
    declared = set(re.findall(r"^(?:int|void|const char \*|rt_ctx \*)\s*(rt_[a-z0-9_]+)\(", hdr, re.M))
    assert declared == set(capi.RT_SYMBOLS), declared ^ set(capi.RT_SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.rt_abi_version() == 2


def test_struct_layouts_match_header():
    assert C.sizeof(capi.ObjectDesc) == 64 and C.sizeof(capi.XformDesc) == 96
    assert C.sizeof(capi.MaterialDesc) == 32 and C.sizeof(capi.TextureDesc) == 32
    assert C.sizeof(capi.CameraDesc) == 52 and C.sizeof(capi.RenderParams) == 32


def test_no_device_is_an_error_not_a_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RtError, match="no CUDA device"):
        Renderer(0)


def test_product_never_references_the_oracle():
    """The product tree must not include, link or load anything under oracle/."""
    pkg = os.path.join(ROOT, "ray-tracing-in-cuda_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and not re.search(r'#include\s+"[^"]*(oracle|orc_)', text), f
                assert not re.search(r"^\s*(import|from)\s+orc\b", text, re.M), f
    out = os.popen(f"ldd {capi.LIB_PATH}").read()
    assert "oracle" not in out


def test_png_and_ppm_writers(lib, tmp_path):
    from PIL import Image
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    renderer.write_png(str(tmp_path / "x.png"), img)
    renderer.write_ppm(str(tmp_path / "x.ppm"), img)
    assert np.array_equal(np.asarray(Image.open(tmp_path / "x.png")), img)
    assert np.array_equal(np.asarray(Image.open(tmp_path / "x.ppm")), img)
    assert np.array_equal(renderer.read_png(str(tmp_path / "x.png")), img)
    # PPM is the reference's P3 text layout (color.cuh:70-95): header then one "r g b" line per pixel, top row first
    lines = open(tmp_path / "x.ppm").read().split("\n")
    assert lines[:3] == ["P3", "53 37", "255"] and lines[3] == " ".join(map(str, img[0, 0]))
    # reader handles PIL-written RGBA / grey / palette / filtered PNGs
    for mode in ("RGBA", "L", "P"):
        Image.fromarray(img).convert(mode).save(tmp_path / f"m_{mode}.png", optimize=True)
        want = np.asarray(Image.open(tmp_path / f"m_{mode}.png").convert("RGB"))
        assert np.array_equal(renderer.read_png(str(tmp_path / f"m_{mode}.png")), want)
    with pytest.raises(RtError):
        renderer.read_png(str(tmp_path / "missing.png"))


def test_crafted_png_is_an_error_not_a_crash(lib, tmp_path):
    """No C++ exception crosses the C ABI: a PNG whose header declares 2^31 x 2^31 pixels over four bytes of data is rejected."""
    import struct
    import zlib

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xffffffff)
    png = (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", 0x7fffffff, 0x7fffffff, 8, 2, 0, 0, 0)) +
           chunk(b"IDAT", zlib.compress(b"\0\0\0\0")) + chunk(b"IEND", b""))
    (tmp_path / "huge.png").write_bytes(png)
    with pytest.raises(RtError):
        renderer.read_png(str(tmp_path / "huge.png"))


def test_animation_rules_follow_the_reference_scripts():
    import json
    basic = json.load(open(os.path.join(ROOT, "tests", "golden", "basic_scene.json")))
    f0 = animation.dna_frame(basic, 0)
    assert f0 == json.load(open(os.path.join(ROOT, "tests", "golden", "dna_000.json")))   # golden was made from dna.py's rules
    assert len(f0["object"]["data"]) == 90 and len(f0["material"]["data"]) == 90 and len(f0["texture"]["data"]) == 90
    f7 = animation.dna_frame(basic, 7)
    assert f7["object"]["data"][2]["rotate"]["angle"] == 90 + 7 and f7["output_file"].endswith("output_007.png")
    blue2 = json.load(open(os.path.join(ROOT, "tests", "golden", "blue2.json")))
    g = animation.blue2_frame(blue2, 3)      # blue2.py: +2 degrees per frame, applied before the frame is written
    cyl = [o for o in g["object"]["data"] if o["type"] == "cylinder"]
    base = [o for o in blue2["object"]["data"] if o["type"] == "cylinder"]
    assert [c["rotate"]["angle"] for c in cyl] == [b["rotate"]["angle"] + 8 for b in base]
    assert animation.frames_for_rank(20, 3, 8) == [3, 11, 19]


def test_cli_keeps_the_reference_exit_code_without_a_device(lib, tmp_path):
    """parallel_compute -f scene.json: exit 99 on failure like check_cuda() (rtweekend.cuh:43-53); never a CPU render."""
    import subprocess
    import torch
    cli = os.path.join(ROOT, "ray-tracing-in-cuda_b200", "parallel_compute")
    assert os.path.exists(cli)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([cli, "-f", os.path.join(ROOT, "tests", "golden", "sample_scene.json")], cwd=tmp_path, capture_output=True, text=True)
    assert p.returncode == 99 and "no CUDA device" in p.stderr
    assert not (tmp_path / "main.ppm").exists()
    p = subprocess.run([cli, "--bogus"], cwd=tmp_path, capture_output=True, text=True)
    assert p.returncode == 99


def test_bench_clock_sampler_degrades_without_a_gpu():
    """bench.py samples clocks through NVML (nvidia-smi as fallback); without a GPU it must report that, not raise."""
    import importlib.util
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    s = bench.ClockSampler(0)
    s.start()
    time.sleep(0.1)
    out = s.stop()
    assert "reasons" in out and "sm_mhz" in out


def test_moving_light_rule_frame_deltas_and_frame_sequence_assembly(lib, tmp_path):
    """README.md:14-17 of the reference (to-do: moving light, video from frames): the moving-light rule turns the emissive spheres
    about the look-at axis and touches nothing else; scene_delta finds the table ranges rt_scene_update needs; write_apng assembles
    frames in order into one animated PNG that an independent reader (PIL) plays back."""
    import json
    import math
    from PIL import Image
    ext = json.load(open(os.path.join(ROOT, "tests", "golden", "sample_scene_ext.json")))
    f0, f30 = animation.moving_light_frame(ext, 0, 120), animation.moving_light_frame(ext, 30, 120)
    mats = ext["material"]["data"]
    moved = 0
    for a, b, c in zip(ext["object"]["data"], f0["object"]["data"], f30["object"]["data"]):
        if a["type"] == "sphere" and mats[a["material"]]["type"] == "diffuse_light":
            cx, _, cz = ext["camera"]["lookat"]
            r = lambda o: math.hypot(o["center"][0] - cx, o["center"][2] - cz)
            assert a["center"] == pytest.approx(b["center"]) and r(c) == pytest.approx(r(a)) and c["center"][1] == a["center"][1]
            assert c["center"] != pytest.approx(a["center"])       # a quarter turn
            moved += 1
        else:
            assert a == b == c
    assert moved == 1 and animation.scene_delta(f0, f30) is None      # triangles / meshes in the scene: full upload
    blue2 = json.load(open(os.path.join(ROOT, "tests", "golden", "blue2.json")))
    a, b = animation.blue2_frame(blue2, 0), animation.blue2_frame(blue2, 1)
    assert animation.scene_delta(a, b) == (2, 6, 0, 4) and animation.scene_delta(a, a) == (0, 0, 0, 0)
    b2 = json.loads(json.dumps(b)); b2["max_depth"] = 7
    assert animation.scene_delta(a, b2) is None
    paths = []
    for k in range(4):
        img = np.zeros((12, 16, 3), np.uint8); img[..., 0] = 60 * k; img[k, k] = 255
        renderer.write_png(str(tmp_path / f"{k:03d}.png"), img); paths.append(str(tmp_path / f"{k:03d}.png"))
    assert animation.write_apng(paths, str(tmp_path / "anim.png"), fps=12) == 4
    im = Image.open(tmp_path / "anim.png")
    assert im.n_frames == 4
    for k in range(4):
        im.seek(k)
        assert np.array_equal(np.asarray(im.convert("RGB")), np.asarray(Image.open(paths[k])))
    assert animation.write_frame_list(paths, str(tmp_path / "frames.txt"), fps=12) == 4
    assert open(tmp_path / "frames.txt").read().count("duration 0.083333") == 4
