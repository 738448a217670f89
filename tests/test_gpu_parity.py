"""GPU parity tests (-m gpu): the CUDA wavefront renderer, called through the C ABI (librt_b200.so), against the
CPU oracle on the same seeded inputs, and against the golden vectors made from the real reference.

Bars: integer work (LBVH keys / order / topology, ray counts) bit-exact; float radiance within 1e-5 per sample
(the north star's deterministic-path tolerance; in practice the images are bit-identical because both sides follow
the same fma contract); converged renders within Monte-Carlo noise of the reference's own images."""
import json
import os

import numpy as np
import pytest

import orc
from rt_b200 import Renderer, RtError, capi, renderer

pytestmark = pytest.mark.gpu
SCENES = ["sample_scene", "blue", "blue2", "dna_000", "rtiow_486"]
G = orc.GOLDEN


@pytest.fixture(scope="module")
def rt():
    r = Renderer(0)
    yield r
    r.close()


def scene_path(name):
    return f"{G}/{name}.json"


@pytest.mark.parametrize("name", SCENES + ["basic_scene", "sample_scene_ext"])
def test_lbvh_bit_exact(rt, name):
    """GPU LBVH build vs the host twin: Morton keys, sorted order, Karras child/parent links bit-exact; boxes equal."""
    rt.upload_json(scene_path(name))
    got = rt.bvh()
    want = orc.lbvh(orc.load_scene(scene_path(name)))
    n = got["n"]
    if n == 0:
        return
    assert (got["n_tree"], got["n_giants"], got["depth"]) == (want["n_tree"], want["n_giants"], want["depth"])
    assert got["n_giants"] == {"rtiow_486": 1, "sample_scene": 1, "sample_scene_ext": 1, "blue": 2, "blue2": 2}.get(name, 0)
    assert np.array_equal(got["morton"], want["morton"])
    assert np.array_equal(got["prim"], want["prim"])
    assert np.array_equal(got["scene_aabb"], want["scene_aabb"])
    assert np.array_equal(got["leaf_aabb"], want["leaf_aabb"])
    if got["n_tree"] >= 2:
        for k in ("left", "right", "parent"):
            assert np.array_equal(got[k], want[k]), k
        assert np.array_equal(got["node_aabb"], want["node_aabb"])


def random_scene(rng, n, with_ext=True):
    """Seeded scene mixing every primitive type (incl. duplicates => equal Morton codes and exact-tie hits)."""
    objs, mats, texs = [], [], []
    texs.append({"type": "solid_color", "color": [0.8, 0.3, 0.3]})
    texs.append({"type": "solid_color", "color": [4, 4, 4]})
    texs.append({"type": "solid_color", "color": [0.2, 0.3, 0.1]})
    texs.append({"type": "checker", "even": 0, "odd": 2})
    mats += [{"type": "lambertian", "texture": 0}, {"type": "lambertian", "texture": 3},
             {"type": "metal", "albedo": [0.7, 0.6, 0.5], "fuzz": 0.2}, {"type": "dielectric", "index_of_refraction": 1.5},
             {"type": "diffuse_light", "texture": 1}]
    kinds = ["sphere", "xy_rect", "xz_rect", "yz_rect", "cylinder"] + (["box", "triangle"] if with_ext else [])
    for i in range(n):
        k = kinds[int(rng.integers(len(kinds)))]
        c = rng.uniform(-4, 4, 3).round(3).tolist()
        m = int(rng.integers(len(mats)))
        if k == "sphere":
            o = {"type": k, "center": c, "radius": float(np.round(rng.uniform(0.2, 0.9), 3)) * (-1 if rng.random() < 0.1 else 1)}
        elif k.endswith("rect"):
            a, b = k[0], k[1]
            o = {"type": k, a + "0": c[0] - 1, a + "1": c[0] + 1, b + "0": c[1] - 0.7, b + "1": c[1] + 0.7, "k": c[2]}
        elif k == "cylinder":
            o = {"type": k, "radius": 0.4, "zmin": -0.8, "zmax": 0.8, "translate": c,
                 "rotate": {"axis": rng.normal(size=3).round(3).tolist(), "angle": float(rng.uniform(0, 360))}}
        elif k == "box":
            o = {"type": k, "min": [-0.5, -0.4, -0.3], "max": [0.5, 0.4, 0.3], "translate": c}
            if rng.random() < 0.7:
                o["rotate"] = {"axis": [0, 1, 0], "angle": float(rng.uniform(0, 90))}
        else:
            v = (np.array(c) + rng.normal(size=(3, 3))).round(3)
            o = {"type": k, "v0": v[0].tolist(), "v1": v[1].tolist(), "v2": v[2].tolist()}
        o["material"] = m
        objs.append(o)
        if rng.random() < 0.08:           # exact duplicate with another material: equal t => "last index wins"
            d = dict(o); d["material"] = int(rng.integers(len(mats))); objs.append(d)
    return {"background": [0.5, 0.7, 0.99], "max_depth": 12, "samples_per_pixel": 4, "width": 80, "height": 45,
            "camera": {"lookfrom": [0, 2, 14], "lookat": [0, 0, 0], "vup": [0, 1, 0], "vfov": 40, "aperture": 0.1},
            "object": {"data": objs}, "material": {"data": mats}, "texture": {"data": texs}}


@pytest.mark.parametrize("n,seed", [(1, 1), (2, 2), (3, 3), (40, 4), (700, 5), (5000, 6)])
def test_lbvh_bit_exact_random(rt, n, seed):
    sc = random_scene(np.random.default_rng(seed), n)
    rt.upload_json_text(json.dumps(sc))
    got = rt.bvh()
    want = orc.lbvh(orc.load_scene(sc))
    assert got["n"] == len(sc["object"]["data"])
    assert (got["n_tree"], got["n_giants"], got["depth"]) == (want["n_tree"], want["n_giants"], want["depth"])
    for k in ("morton", "prim", "leaf_aabb", "scene_aabb"):
        assert np.array_equal(got[k], want[k]), k
    if got["n_tree"] >= 2:
        for k in ("left", "right", "parent", "node_aabb"):
            assert np.array_equal(got[k], want[k]), k


@pytest.mark.parametrize("name", SCENES)
def test_aov_matches_oracle_and_reference(rt, name):
    """Deterministic path (pixel-centre primary rays, first hit): CUDA == oracle on every field, and the colour channels
    are within 1e-5 of the REAL reference's output (golden)."""
    ref = np.load(f"{G}/aov_{name}.npz")["aov"]
    H, W, _ = ref.shape
    rt.upload_json(scene_path(name))
    rt.override(width=W, height=H)
    got = rt.render_aov()
    want = orc.aov(orc.load_scene(scene_path(name), width=W, height=H))
    assert np.array_equal(got[..., 0], want[..., 0]) and np.array_equal(got[..., 11], want[..., 11]), "BVH traversal != linear walk"
    assert np.abs(got - want).max() <= 1e-5 * max(1.0, np.abs(want).max())
    same = (got[..., 0] == ref[..., 0]) & (got[..., 11] == ref[..., 11])
    assert (~same).sum() <= (5 if name == "rtiow_486" else 0)
    assert np.abs(got[..., 13:22] - ref[..., 13:22])[same].max() <= 1e-5


def constant_image_sum(bg, spp):
    """spp x bg as the fixed-order accumulation adds it: sample by sample within each of the 8 sample chunks (main.cu:95-104),
    chunk sums combined in chunk order (rt.h: rt_sample_chunk)."""
    from rt_b200 import multi_gpu
    total = None
    for k in range(8):
        b, e = multi_gpu.sample_chunk(spp, k)
        if e <= b:
            continue
        c = np.zeros(3, np.float32)
        for _ in range(e - b):
            c = c + bg
        total = c if total is None else total + c
    return total


def scene_dict(name, sampler=None):
    with open(scene_path(name)) as f:
        d = json.load(f)
    if sampler:
        d["sampler"] = sampler        # "rejection": the reference's loops instead of the direct draws (RT_SAMPLING_DIRECT)
    return d


@pytest.mark.parametrize("name,spp,sampler", [("sample_scene", 8, None), ("blue", 4, None), ("blue2", 8, None), ("dna_000", 4, None),
                                              ("rtiow_486", 8, None), ("basic_scene", 3, None), ("sample_scene_ext", 6, None),
                                              ("rtiow_486", 6, "rejection"), ("blue2", 6, "rejection"), ("sample_scene_ext", 4, "rejection")])
def test_render_matches_oracle_sample_for_sample(rt, name, spp, sampler):
    """Full path tracing, same Philox streams: per-pixel radiance sums within 1e-5 per sample of the brute-force oracle,
    and exactly the same number of path segments — with the direct samplers (default) and with the reference's rejection loops."""
    W, H = 96, 54
    d = scene_dict(name, sampler)
    rt.upload_json_text(json.dumps(d), base_dir=G)
    rt.override(width=W, height=H, spp=spp)
    rt.stats_reset()
    rt.render(seed=0x2023)
    got, _, st = rt.download()
    want, nrays = orc.render(orc.load_scene(d, width=W, height=H, spp=spp, base_dir=G), seed=0x2023)
    assert st["rays"] == nrays
    err = np.abs(got - want)
    bad = (err > 1e-5 * spp).any(-1).sum()
    assert bad == 0, f"{bad} pixels differ, max {err.max()}"


@pytest.mark.parametrize("name,spp", [("rtiow_486", 3), ("blue2", 4), ("sample_scene_ext", 3), ("dna_000", 2)])
def test_traversal_counters_match_the_host_twin(rt, name, spp):
    """The device's event counters (inner-node visits, box tests, leaf tests per primitive type) against a serial count of the
    same walk (oracle/orc_trav.c: giants first, ordered descent, far child pushed with its entry distance): integer work, exact.
    Bounce 0 goes through the tree too (RT_RENDER_NO_TILE_LISTS), like every ray of the host twin."""
    W, H = 96, 54
    rt.upload_json(scene_path(name))
    rt.override(width=W, height=H, spp=spp)
    rt.stats_reset()
    rt.render(seed=0x51, flags=capi.RT_RENDER_COUNT_EVENTS | capi.RT_RENDER_NO_TILE_LISTS)
    got, _, st = rt.download()
    want, nrays, cnt = orc.render_bvh(orc.load_scene(scene_path(name), width=W, height=H, spp=spp), seed=0x51)
    assert st["rays"] == nrays and np.array_equal(got, want)
    for k in ("n_node_visits", "n_box_tests", "n_sphere_reject", "n_sphere_accept", "n_rect_tests", "n_cyl_reject", "n_cyl_accept",
              "n_tri_tests", "n_boxprim_tests"):
        assert st[k] == cnt[k], (k, st[k], cnt[k])


def test_render_options_do_not_change_the_image(rt):
    """Material sorting, pass size, split sample ranges and event counting change scheduling only: images are bit-identical."""
    W, H, spp = 64, 36, 6
    rt.upload_json(scene_path("blue2"))
    rt.override(width=W, height=H, spp=spp)
    rt.render(seed=5)
    base, _, _ = rt.download()
    rt.render(seed=5, flags=capi.RT_RENDER_SORT)
    assert np.array_equal(rt.download()[0], base)
    rt.render(seed=5, max_paths_per_pass=W * H * 2)
    assert np.array_equal(rt.download()[0], base)
    rt.render(seed=5, flags=capi.RT_RENDER_SERIAL)                       # one stream instead of two passes in flight
    assert np.array_equal(rt.download()[0], base)
    rt.stats_reset(); rt.render(seed=5, flags=capi.RT_RENDER_NO_TAIL)     # queue schedule to the last bounce instead of the tail launch
    img, _, st_q = rt.download()
    rt.stats_reset(); rt.render(seed=5)
    _, _, st_t = rt.download()
    assert np.array_equal(img, base) and st_q["rays"] == st_t["rays"] and st_q["launches"] > st_t["launches"]
    rt.render(seed=5, flags=capi.RT_RENDER_NO_TILE_LISTS)                # bounce 0 through the tree instead of the beam lists
    assert np.array_equal(rt.download()[0], base)
    rt.render(seed=5, flags=capi.RT_RENDER_SERIAL, max_paths_per_pass=W * H)
    assert np.array_equal(rt.download()[0], base)
    rt.render(seed=5, max_paths_per_pass=W * H)                           # six one-sample passes alternating between the streams
    assert np.array_equal(rt.download()[0], base)
    rt.render(seed=5, flags=capi.RT_RENDER_COUNT_EVENTS)
    img, _, st = rt.download()
    assert np.array_equal(img, base) and st["n_box_tests"] > 0 and st["n_shade"] > 0
    rt.render(seed=5, spp_begin=0, spp_count=4)
    rt.render(seed=5, spp_begin=4, spp_count=2, flags=capi.RT_RENDER_KEEP_ACCUM)
    assert np.array_equal(rt.download()[0], base)
    rt.render(seed=6)
    assert not np.array_equal(rt.download()[0], base)


# RT_SOAK=n: n more seeded scenes (a wider sweep for one-off runs after traversal changes)
@pytest.mark.parametrize("seed", [11, 12] + list(range(100, 100 + int(os.environ.get("RT_SOAK", "0")))))
def test_random_scenes_all_primitives(rt, seed):
    """Every primitive type incl. box/triangle extensions, duplicates (exact ties) and negative radii vs the oracle."""
    sc = random_scene(np.random.default_rng(seed), 60)
    rt.upload_json_text(json.dumps(sc))
    got_aov = rt.render_aov()
    osc = orc.load_scene(sc)
    want_aov = orc.aov(osc)
    assert np.array_equal(got_aov[..., 11], want_aov[..., 11]), "closest-hit object differs (tie rule?)"
    assert np.abs(got_aov - want_aov).max() <= 1e-5 * max(1.0, np.abs(want_aov).max())
    rt.stats_reset()
    rt.render(seed=3)
    got, _, st = rt.download()
    want, nrays = orc.render(osc, seed=3)
    assert st["rays"] == nrays
    assert np.abs(got - want).max() <= 1e-5 * sc["samples_per_pixel"]


def test_json_reader_equals_table_upload(rt):
    """rt_scene_upload_json (C++ reader) and rt_scene_upload (tables from the independent Python loader) agree bit for bit."""
    for name in ("sample_scene", "blue2"):
        rt.upload_json(scene_path(name)); rt.override(width=64, height=36, spp=2); rt.render(seed=9)
        a = rt.download()[0]
        rt.upload(orc.load_scene(scene_path(name), width=64, height=36, spp=2)); rt.render(seed=9)
        b = rt.download()[0]
        assert np.array_equal(a, b)


@pytest.mark.parametrize("name", ["sample_scene", "blue2", "dna_000", "rtiow_486"])
def test_converged_against_reference(rt, name):
    """Converged render at the golden's spp vs the REAL reference's two independent converged images:
    RMSE <= 1.5 x the reference's own run-to-run RMSE; PSNR reported."""
    g = np.load(f"{G}/conv_{name}.npz")
    a, b, spp = g["a"], g["b"], int(g["spp"])
    H, W, _ = a.shape
    rt.upload_json(scene_path(name)); rt.override(width=W, height=H, spp=spp)
    rt.render(seed=77)
    img = rt.download()[0] / spp

    def rmse(x, y):
        return float(np.sqrt(np.mean((x - y) ** 2)))
    noise = rmse(a, b)
    assert rmse(img, a) < 1.5 * noise and rmse(img, b) < 1.5 * noise
    ref_mean = 0.5 * (a + b)
    psnr = 10 * np.log10(1.0 / max(np.mean((np.clip(img, 0, 1) - np.clip(ref_mean, 0, 1)) ** 2), 1e-12))
    print(f"{name}: rmse={rmse(img, ref_mean):.5f} noise={noise:.5f} psnr={psnr:.1f} dB")
    assert psnr > 30.0


def test_full_size_properties(rt):
    """BASELINE-size properties that need no oracle: the literal basic_scene.json (0 objects) at 1920x1080 must give
    exactly spp * background in every pixel; sample-range linearity holds at full size on blue2."""
    rt.upload_json(scene_path("basic_scene")); rt.override(width=1920, height=1080, spp=16)
    rt.stats_reset(); rt.render(seed=1)
    img, _, st = rt.download()
    bg = np.array([0.0005, 0.0007, 0.00099], np.float32)
    want = constant_image_sum(bg, 16)
    assert np.array_equal(img, np.broadcast_to(want, img.shape)) and st["rays"] == 1920 * 1080 * 16
    rt.upload_json(scene_path("blue2")); rt.override(spp=4)
    rt.render(seed=2)
    full = rt.download()[0]
    rt.render(seed=2, spp_begin=0, spp_count=2); rt.render(seed=2, spp_begin=2, spp_count=2, flags=capi.RT_RENDER_KEEP_ACCUM)
    assert np.array_equal(rt.download()[0], full)
    assert np.isfinite(full).all() and full.min() >= 0


def test_tonemap_and_image_writers(rt, tmp_path):
    rt.upload_json(scene_path("sample_scene")); rt.override(width=64, height=36, spp=4); rt.render(seed=1)
    s, png8, _ = rt.download(gamma=0)
    _, ppm8, _ = rt.download(gamma=1)
    assert np.array_equal(png8, orc.tonemap(s, 4, 0)) and np.array_equal(ppm8, orc.tonemap(s, 4, 1))
    renderer.write_png(str(tmp_path / "a.png"), png8); renderer.write_ppm(str(tmp_path / "a.ppm"), ppm8)
    assert np.array_equal(renderer.read_png(str(tmp_path / "a.png")), png8)
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(tmp_path / "a.png")), png8)
    assert np.array_equal(np.asarray(Image.open(tmp_path / "a.ppm")), ppm8)


def test_errors_are_reported_not_fatal(rt):
    with pytest.raises(RtError, match="cannot open"):
        rt.upload_json("/nonexistent/scene.json")
    with pytest.raises(RtError, match="JSON"):
        rt.upload_json_text("{ not json")
    with pytest.raises(RtError, match="missing material"):
        rt.upload_json_text(json.dumps({"background": [0, 0, 0], "max_depth": 5, "samples_per_pixel": 1, "width": 8, "height": 8,
                                        "camera": {"lookfrom": [0, 0, 1], "lookat": [0, 0, 0], "vup": [0, 1, 0], "vfov": 20, "aperture": 0},
                                        "object": {"data": [{"type": "sphere", "center": [0, 0, 0], "radius": 1, "material": 3}]},
                                        "material": {"data": []}, "texture": {"data": []}}))
    rt.upload_json(scene_path("sample_scene"))      # the context is still usable
    assert rt.info().n_objects == 6


def test_cli_matches_the_library_and_the_reference_contract(rt, tmp_path):
    """parallel_compute -f scene.json (main.cu:448-522): writes main.ppm + output_file, exit 0; pixels equal rt_download's."""
    import subprocess
    from PIL import Image
    d = json.load(open(scene_path("sample_scene")))
    d.update(width=64, height=36, samples_per_pixel=3, output_file=str(tmp_path / "out.png"))
    f = tmp_path / "s.json"
    f.write_text(json.dumps(d))
    cli = os.path.join(os.path.dirname(capi.LIB_PATH), "parallel_compute")
    p = subprocess.run([cli, "-f", str(f)], cwd=tmp_path, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    rt.upload_json(str(f))
    rt.render(seed=0x2023)
    _, png8, _ = rt.download(gamma=0)
    _, ppm8, _ = rt.download(gamma=1, want_sum=False)
    assert np.array_equal(np.asarray(Image.open(tmp_path / "out.png")), png8)
    assert np.array_equal(np.asarray(Image.open(tmp_path / "main.ppm")), ppm8)
    bad = tmp_path / "bad.json"
    bad.write_text("{ not json")
    assert subprocess.run([cli, "-f", str(bad)], cwd=tmp_path, capture_output=True).returncode == 99


def test_triangle_mesh_tables_match_the_oracle(rt):
    """BASELINE config 4 at test size: a seeded displaced-grid mesh handed over as flat tables (rt_scene_upload);
    LBVH bit-exact vs the host twin, render bit-identical to the brute-force oracle (parity unpinned by reference:
    the reference has no C++ triangle code)."""
    from rt_b200 import scenes
    ts = scenes.mesh_scene(12, 9, width=64, height=36, spp=3, seed=5, max_depth=8)
    rt.upload(ts)
    got = rt.bvh(); want = orc.lbvh(ts)
    for k in ("morton", "prim", "left", "right", "parent", "node_aabb", "leaf_aabb"):
        assert np.array_equal(got[k], want[k]), k
    rt.stats_reset()
    rt.render(seed=3, flags=capi.RT_RENDER_COUNT_EVENTS)
    img, _, st = rt.download()
    ref, nrays = orc.render(ts, seed=3)
    assert st["rays"] == nrays and st["n_tri_tests"] > 0
    assert np.abs(img - ref).max() <= 1e-5 * 3


def test_million_triangle_mesh_builds_and_renders(rt):
    """Full-size config 4 scene (1,002,528 triangles + 4 other objects): the LBVH builds on the device, the traversal
    stack does not overflow (COUNT_EVENTS turns an overflow into an error), every path terminates, the image is finite,
    and a second render of the same seed is bit-identical (fixed-order accumulation, counter-based RNG)."""
    from rt_b200 import scenes
    ts = scenes.mesh_scene(708, 708, width=160, height=90, spp=2)
    info = rt.upload(ts)
    assert info.n_objects == 2 * 708 * 708 + 4
    b = rt.bvh()
    assert np.all(np.diff(b["morton"].astype(np.uint64)) >= 0) and sorted(b["prim"].tolist()) == list(range(info.n_objects))
    rt.stats_reset()
    rt.render(seed=1, flags=capi.RT_RENDER_COUNT_EVENTS)
    a, _, st = rt.download()
    assert np.isfinite(a).all() and st["rays"] >= 160 * 90 * 2 and st["n_tri_tests"] > 0
    rt.render(seed=1)
    a2, _, _ = rt.download()
    assert np.array_equal(a, a2)


def test_animation_driver_frames_resume_and_files(rt, tmp_path):
    """blue2.py's rule in-process: every frame's PNG equals a stand-alone render of that frame's JSON; a second run
    skips the frames already on disk (frame-level resume)."""
    from PIL import Image
    from rt_b200 import animation
    base = json.load(open(scene_path("blue2")))
    res = animation.render_animation(rt, base, animation.blue2_frame, [0, 3, 5], str(tmp_path), width=64, height=36, spp=2)
    assert res["frames"] == 3 and res["skipped"] == 0 and res["samples"] == 3 * 64 * 36 * 2 and res["launches"] > 0
    for f in (0, 3, 5):
        d = animation.blue2_frame(base, f)
        d["camera"]["aspect_ratio"] = d["width"] / d["height"]
        d.update(width=64, height=36, samples_per_pixel=2)
        rt.upload_json_text(json.dumps(d))
        rt.render(seed=0x2023)
        _, png8, _ = rt.download(gamma=0)
        assert np.array_equal(np.asarray(Image.open(tmp_path / f"{f:03d}.png")), png8)
    again = animation.render_animation(rt, base, animation.blue2_frame, [0, 3, 5, 6], str(tmp_path), width=64, height=36, spp=2)
    assert again["frames"] == 1 and again["skipped"] == 3


@pytest.mark.parametrize("W,H,spp", [(67, 35, 3), (2, 2, 1), (9, 5, 7), (130, 3, 2)])
def test_odd_image_sizes(rt, W, H, spp):
    """Sizes that are not multiples of the 8x4 camera tile, of the shade chunk or of a 16-byte TMA copy (odd tails):
    image and ray count still equal the oracle's, with and without the in-CTA material sort."""
    rt.upload_json(scene_path("blue2"))
    rt.override(width=W, height=H, spp=spp)
    want, nrays = orc.render(orc.load_scene(scene_path("blue2"), width=W, height=H, spp=spp), seed=9)
    for flags in (0, capi.RT_RENDER_SORT):
        rt.stats_reset()
        rt.render(seed=9, flags=flags)
        got, _, st = rt.download()
        assert st["rays"] == nrays
        assert np.abs(got - want).max() <= 1e-5 * spp


def _nee_scene(extra_lights=True):
    with open(scene_path("sample_scene_ext")) as f:
        d = json.load(f)
    d["sampling"] = "reference_cpu"
    if extra_lights:     # a triangle light and an emissive box (a face is picked by area)
        light = next(i for i, m in enumerate(d["material"]["data"]) if m["type"] == "diffuse_light")
        d["object"]["data"].append({"type": "triangle", "v0": [-3.0, 2.5, -2.0], "v1": [-2.0, 2.5, -2.0], "v2": [-2.5, 2.9, -1.2], "material": light})
        d["object"]["data"].append({"type": "box", "min": [2.6, -0.4, 1.0], "max": [2.9, -0.1, 1.3], "material": light})
        # sphere and rect lights that carry a rotate / translate instance are sampled in world space
        d["object"]["data"].append({"type": "sphere", "center": [0.1, 0.0, 0.05], "radius": 0.25, "material": light,
                                    "rotate": {"axis": [0.2, 1, 0.1], "angle": 33}, "translate": [-1.2, 1.6, 0.8]})
        d["object"]["data"].append({"type": "xy_rect", "x0": -0.3, "x1": 0.3, "y0": -0.2, "y1": 0.2, "k": 0.1, "material": light,
                                    "rotate": {"axis": [1, 0.3, 0], "angle": -50}, "translate": [1.4, 1.9, 0.2]})
        # a degenerate emitter (zero area) is left out of the light list: nothing to sample, no 1/area in the MIS weight
        d["object"]["data"].append({"type": "triangle", "v0": [0.5, 2.0, 0.5], "v1": [0.6, 2.0, 0.5], "v2": [0.7, 2.0, 0.5], "material": light})
    return d


@pytest.mark.parametrize("extra,depth,mis", [(False, 0, False), (True, 0, False), (True, 3, False), (True, 0, True), (True, 3, True)])
def test_nee_matches_oracle_sample_for_sample(rt, extra, depth, mis):
    """RT_RENDER_NEE (the shadow stage): light sampling at every lambertian vertex + any-hit occlusion rays, same
    Philox streams as oracle/orc_nee.c: radiance within 1e-5 per sample, equal path-segment and shadow-ray counts."""
    W, H, spp = 96, 54, 5
    nee_flag = capi.RT_RENDER_NEE_MIS if mis else capi.RT_RENDER_NEE        # MIS: balance-heuristic weights (oracle/orc_nee.c)
    d = _nee_scene(extra)
    rt.upload_json_text(json.dumps(d), base_dir=G)
    rt.override(width=W, height=H, spp=spp, max_depth=depth)
    rt.stats_reset()
    rt.render(seed=0x77, flags=nee_flag)
    got, _, st = rt.download()
    want, nrays, nshadow = orc.render_ex(orc.load_scene(d, width=W, height=H, spp=spp, max_depth=depth or None, base_dir=G), seed=0x77,
                                         flags=nee_flag)
    assert nshadow > 0 and st["shadow_rays"] == nshadow and st["rays"] == nrays
    err = np.abs(got - want)
    bad = (err > 1e-5 * spp * max(1.0, want.max() / spp)).any(-1).sum()
    assert bad == 0, f"{bad} pixels differ, max {err.max()}"
    # scheduling options leave the NEE image unchanged too
    rt.render(seed=0x77, flags=nee_flag | capi.RT_RENDER_SORT, max_paths_per_pass=W * H * 2)
    assert np.array_equal(rt.download()[0], got)


def test_nee_same_expectation_lower_variance(rt):
    """Light sampling (eight lights of six kinds, three of them rotated instances) and light sampling + MIS converge to the plain
    estimator's image, and with MIS the image is the less noisy one.  Light sampling ALONE is not on this scene: the emissive box stands
    0.1 above the floor, and the 1/d^2 of a point sampled next to the shaded point is exactly what the balance heuristic is there for
    (oracle, same seeds: plain 0.0068, light sampling 0.0110, MIS 0.0042 mean squared difference of two renders)."""
    W, H, spp = 64, 36, 256
    d = _nee_scene(True)
    d["background"] = [0.02, 0.02, 0.03]
    rt.upload_json_text(json.dumps(d), base_dir=G)
    rt.override(width=W, height=H, spp=spp)
    rt.render(seed=1); plain = rt.download()[0] / spp
    rt.render(seed=1, flags=capi.RT_RENDER_NEE); nee = rt.download()[0] / spp
    assert abs(nee.mean() - plain.mean()) <= 0.03 * plain.mean()
    rt.render(seed=1, flags=capi.RT_RENDER_NEE_MIS); mis = rt.download()[0] / spp
    assert abs(mis.mean() - plain.mean()) <= 0.03 * plain.mean()
    rt.render(seed=2, flags=capi.RT_RENDER_NEE_MIS); mis2 = rt.download()[0] / spp
    rt.render(seed=2); plain2 = rt.download()[0] / spp
    assert np.mean((mis - mis2) ** 2) < 0.8 * np.mean((plain - plain2) ** 2)


def test_nee_needs_cpu_sampling_convention(rt):
    rt.upload_json(scene_path("sample_scene_ext"))
    rt.override(width=32, height=18, spp=1)
    with pytest.raises(RtError):
        rt.render(flags=capi.RT_RENDER_NEE)
    rt.render()            # the context stays usable


def test_axis_parallel_rays_and_faces_on_the_ray_plane(rt):
    """Rays with a direction component of exactly 0 (centre column / centre row of an axis-aligned camera at the origin of
    that axis) against geometry whose box faces lie exactly in the ray's plane: the slab pair of the parallel axis must act
    as an inside/outside test (no NaN, nothing culled that the linear walk accepts).  Also a full render for the counts."""
    W, H = 96, 54           # even sizes: pixel column 47 / row 26 have u = v = 0.5 exactly
    tri = lambda a, b, c, m: {"type": "triangle", "v0": a, "v1": b, "v2": c, "material": m}
    objs = []
    for k in range(-3, 3):  # a strip of triangle pairs sharing edges on the plane x = 0 and on the plane y = 0
        z = float(k)
        objs += [tri([0, -1, z], [1, -1, z], [0, 1, z - 0.5], 0), tri([0, -1, z], [0, 1, z - 0.5], [-1, -1, z], 1),
                 tri([-2, 0, z], [2, 0, z], [0, 0, z - 0.9], 2)]
    objs += [{"type": "box", "min": [0, 0, -6], "max": [1.5, 1.5, -5], "material": 0},
             {"type": "box", "min": [-1.5, -1.5, -6], "max": [0, 0, -5], "material": 1},
             {"type": "xy_rect", "x0": -4, "x1": 0, "y0": 0, "y1": 3, "k": -7, "material": 2},
             {"type": "xy_rect", "x0": 0, "x1": 4, "y0": -3, "y1": 0, "k": -7, "material": 0},
             {"type": "sphere", "center": [0, 0, -9], "radius": 1.0, "material": 3},
             {"type": "yz_rect", "y0": -3, "y1": 3, "z0": -12, "z1": 4, "k": 0.0, "material": 1},      # edge-on in the plane x = 0
             {"type": "xz_rect", "x0": -3, "x1": 3, "z0": -12, "z1": 4, "k": 0.0, "material": 2}]      # edge-on in the plane y = 0
    d = {"output_file": "t.png", "background": [0.6, 0.7, 0.9], "max_depth": 8, "samples_per_pixel": 4, "width": W, "height": H,
         "camera": {"lookfrom": [0, 0, 5], "lookat": [0, 0, 0], "vup": [0, 1, 0], "vfov": 40, "aperture": 0.0},
         "object": {"data": objs},
         "material": {"data": [{"type": "lambertian", "texture": 0}, {"type": "lambertian", "texture": 1},
                               {"type": "metal", "albedo": [0.8, 0.8, 0.8], "fuzz": 0.1}, {"type": "dielectric", "index_of_refraction": 1.5}]},
         "texture": {"data": [{"type": "solid_color", "color": [0.8, 0.3, 0.3]}, {"type": "solid_color", "color": [0.3, 0.8, 0.3]}]}}
    rt.upload_json_text(json.dumps(d))
    got = rt.render_aov()
    osc = orc.load_scene(d)
    want = orc.aov(osc)
    # rays that lie IN the plane of an edge-on rect get t = 0/0 = NaN from the reference's rect test (object.cuh:105-110
    # accepts it: NaN fails both range comparisons); the one ray in BOTH planes has two NaN hits and no defined winner
    ok = ~np.isnan(want[..., 1])
    assert np.array_equal(np.isnan(got[..., 1]), ~ok)
    assert np.array_equal(got[..., 0], want[..., 0]) and np.array_equal(got[..., 11][ok], want[..., 11][ok])
    assert (got[..., 11] != want[..., 11]).sum() <= 1
    assert np.abs(got - want)[ok].max() <= 1e-5 * max(1.0, np.abs(want[ok]).max())
    assert want[:, 47, 0].sum() > 0 and want[26, :, 0].sum() > 0          # the degenerate column / row do hit things
    rt.stats_reset()
    rt.render(seed=9)
    img, _, st = rt.download()
    ref, nrays = orc.render(osc, seed=9)
    assert st["rays"] == nrays
    assert (np.abs(img - ref) > 1e-5 * 4).any(-1).sum() == 0


def test_download_into_pinned_buffers(rt):
    """rt_host_alloc targets: same bytes as a download into ordinary numpy arrays; buffers are reusable across renders."""
    rt.upload_json(scene_path("blue2"))
    rt.override(width=64, height=36, spp=3)
    rt.render(seed=4)
    a_sum, a_rgb, _ = rt.download(gamma=1)
    p_sum, p_rgb = rt.pinned_empty((36, 64, 3), np.float32), rt.pinned_empty((36, 64, 3), np.uint8)
    b_sum, b_rgb, _ = rt.download(gamma=1, out_sum=p_sum, out_rgb8=p_rgb)
    assert b_sum is p_sum and b_rgb is p_rgb
    assert np.array_equal(a_sum, p_sum) and np.array_equal(a_rgb, p_rgb)
    rt.render(seed=5)
    rt.download(gamma=1, out_sum=p_sum, out_rgb8=p_rgb)
    assert not np.array_equal(a_sum, p_sum)


@pytest.mark.parametrize("name,W,H", [("rtiow_486", 256, 144), ("rtiow_486", 200, 120), ("sample_scene_ext", 64, 36), ("blue2", 72, 44), ("dna_000", 64, 36)])
def test_primary_hits_from_tile_beams(rt, name, W, H):
    """Bounce-0 hits resolved in the generate stage from the per-tile candidate lists (thin-lens beams for rtiow_486):
    identical to full traversal (flag) and to the brute-force oracle, on fully tiled images (at
    low resolutions some beams exceed RT_TILE_K candidates and go through the fix-up launch, which is covered as well)."""
    spp = 6
    rt.upload_json(scene_path(name))
    rt.override(width=W, height=H, spp=spp)
    rt.stats_reset()
    rt.render(seed=21, flags=capi.RT_RENDER_COUNT_EVENTS)
    got, _, st = rt.download()
    rt.stats_reset()
    rt.render(seed=21, flags=capi.RT_RENDER_COUNT_EVENTS | capi.RT_RENDER_NO_TILE_LISTS)
    full, _, st2 = rt.download()
    assert np.array_equal(got, full) and st["rays"] == st2["rays"]
    assert st["n_box_tests"] < st2["n_box_tests"]        # bounce 0 really skipped the tree (where its tile has a list)
    rt.render(seed=21)                                   # the default schedule (two pass streams) gives the same image
    assert np.array_equal(rt.download()[0], got)
    want, nrays = orc.render(orc.load_scene(scene_path(name), width=W, height=H, spp=spp), seed=21)
    assert st["rays"] == nrays
    assert (np.abs(got - want) > 1e-5 * spp).any(-1).sum() == 0


@pytest.mark.parametrize("name,W,H,spp", [("basic_scene", 128, 72, 5), ("dna_000", 256, 144, 4), ("rtiow_486", 256, 144, 4),
                                          ("rtiow_486", 16, 12, 40), ("blue2", 40, 28, 9), ("dna_000", 200, 100, 3)])
def test_empty_tiles_skip_the_queues(rt, name, W, H, spp):
    """Tiles whose camera beam touches no leaf box (background only) never enter the queues: generate writes at most their
    radiance, shade skips their chunks, accumulate adds a constant background sample by sample.  Same image bit for bit
    and same ray count as the plain schedule (every path through the tree), small passes included (chunks that span
    several sample planes, images smaller than a chunk, partial tiles)."""
    rt.upload_json(scene_path(name))
    rt.override(width=W, height=H, spp=spp)
    rt.stats_reset()
    rt.render(seed=33, flags=capi.RT_RENDER_NO_TILE_LISTS)
    full, _, st_full = rt.download()
    for kw in ({}, {"max_paths_per_pass": W * H * 2}, {"flags": capi.RT_RENDER_SERIAL, "max_paths_per_pass": W * H}):
        rt.stats_reset()
        rt.render(seed=33, **kw)
        got, _, st = rt.download()
        assert np.array_equal(got, full), kw
        assert st["rays"] == st_full["rays"], kw
    if name == "basic_scene":       # no objects: every pixel is spp x background, added one sample at a time (main.cu:95-104)
        bg = np.array(json.load(open(scene_path(name)))["background"], np.float32)
        assert np.array_equal(full, np.broadcast_to(constant_image_sum(bg, spp), full.shape))
        assert st_full["rays"] == W * H * spp


# ---------------------------------------------------------------------------------------------------------------------------
# parity at the sizes that are timed (bench.py), fixed-order accumulation, several GPUs
# ---------------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["rtiow_486", "blue2"])
def test_full_hd_render_matches_the_oracle_pixel_for_pixel(rt, name):
    """The configuration bench.py times — 1920 x 1080, tile_mode 1, passes alternating between the two pass streams — against
    the multi-threaded oracle: every pixel within 1e-5 per sample (in practice bit-identical) and the same ray count.  3 spp = three
    sample chunks = three passes.  rtiow_486 goes through the host twin of the traversal (orc_trav.c, itself checked against the
    linear walk in the CPU tests: the brute-force walk over 486 spheres would take minutes here); blue2 through the linear walk."""
    W, H, spp = 1920, 1080, 3
    d = scene_dict(name)
    d["camera"].setdefault("aspect_ratio", d["width"] / d["height"])
    rt.upload_json_text(json.dumps(d), base_dir=G)
    rt.override(width=W, height=H, spp=spp)
    rt.stats_reset()
    rt.render(seed=0x2023, max_paths_per_pass=W * H)          # one sample plane per pass: three passes over the two pass streams
    got, _, st = rt.download()
    assert st["passes"] >= 3
    sc = orc.load_scene(d, width=W, height=H, spp=spp, base_dir=G)
    want, nrays = (orc.render_bvh(sc, seed=0x2023)[:2] if name == "rtiow_486" else orc.render(sc, seed=0x2023))
    assert st["rays"] == nrays
    err = np.abs(got - want)
    assert (err > 1e-5 * spp).any(-1).sum() == 0, err.max()


def test_4k_mesh_render_matches_the_host_twin_on_a_pixel_subset(rt):
    """BASELINE config 4 at its image size: a displaced-grid triangle mesh at 3840 x 2160, every 97th pixel compared with the
    host twin of the traversal (the tree is read from global memory here: 64-byte nodes, one pass stream)."""
    from rt_b200 import scenes
    ts = scenes.mesh_scene(96, 96, width=3840, height=2160, spp=2, seed=5, max_depth=6)
    rt.upload(ts)
    rt.render(seed=3)
    got, _, _ = rt.download()
    want, _ = orc.render_bvh_strided(ts, 3, 97)
    sel = (np.arange(3840 * 2160) % 97 == 0).reshape(2160, 3840)
    assert np.abs(got[sel] - want[sel]).max() <= 1e-5 * 2
    assert np.abs(want[sel]).max() > 0


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_rank_shards_rendered_one_after_the_other_give_the_one_call_image(rt, world):
    """Fixed-order accumulation: the sample chunks depend on the total sample count only, so the shards of `world` ranks — here
    rendered in turn on one GPU with RT_RENDER_KEEP_ACCUM — fold to exactly the image of one rt_render call."""
    from rt_b200 import multi_gpu
    W, H, spp = 64, 36, 21
    rt.upload_json(scene_path("sample_scene_ext")); rt.override(width=W, height=H, spp=spp)
    rt.render(seed=8)
    full, _, _ = rt.download()
    for r in range(world):
        s0, s1 = multi_gpu.spp_slice(spp, r, world)
        if s1 > s0:
            rt.render(seed=8, spp_begin=s0, spp_count=s1 - s0, total_spp=spp, flags=capi.RT_RENDER_KEEP_ACCUM if r else 0)
    assert np.array_equal(rt.download()[0], full)
    want, _ = orc.render(orc.load_scene(scene_path("sample_scene_ext"), width=W, height=H, spp=spp, base_dir=G), seed=8)
    assert np.array_equal(full, want)                       # the oracle states the same chunk order


def test_device_group_image_is_bit_identical_for_any_gpu_count():
    """rt_group_* (one process, n devices, NCCL behind the C ABI): the image and the ray count for 1, 2, 4, 8 devices (as many as
    the box has) equal the single-context render bit for bit."""
    import torch
    from rt_b200 import renderer as rmod
    W, H, spp = 160, 90, 24
    one = Renderer(0)
    one.upload_json(scene_path("rtiow_486")); one.override(width=W, height=H, spp=spp)
    one.stats_reset(); one.render(seed=5)
    want, _, st1 = one.download()
    one.close()
    for n in (1, 2, 4, 8):
        if n > torch.cuda.device_count():
            break
        g = rmod.Group(list(range(n)))
        g.upload_json(scene_path("rtiow_486")); g.override(width=W, height=H, spp=spp)
        g.render(seed=5)
        got, _, st = g.download()
        assert np.array_equal(got, want), n
        assert st["rays"] == st1["rays"], n
        g.render(seed=6)                                    # the group is reusable
        assert not np.array_equal(g.download()[0], want)
        g.close()


def test_incremental_animation_frames_equal_full_uploads(rt, tmp_path):
    """rt_scene_update (table entries + device rebuild, transforms from rt_make_xform) instead of JSON text per frame: blue2.py's
    rule and a moving emissive sphere give the same PNG bytes as the per-frame JSON path; an invalid update leaves the scene intact."""
    from PIL import Image
    from rt_b200 import animation
    base = json.load(open(scene_path("blue2")))
    full, inc = tmp_path / "full", tmp_path / "inc"
    full.mkdir(); inc.mkdir()
    r1 = animation.render_animation(rt, base, animation.blue2_frame, [0, 1, 2, 5], str(full), width=64, height=36, spp=2)
    r2 = animation.render_animation(rt, base, animation.blue2_frame, [0, 1, 2, 5], str(inc), width=64, height=36, spp=2, incremental=True)
    assert r1["incremental_updates"] == 0 and r2["incremental_updates"] == 3 and r1["rays"] == r2["rays"]
    for f in (0, 1, 2, 5):
        assert np.array_equal(np.asarray(Image.open(full / f"{f:03d}.png")), np.asarray(Image.open(inc / f"{f:03d}.png")))
    lit = json.loads(json.dumps(base))
    lit["texture"]["data"].append({"type": "solid_color", "color": [5, 5, 4]})
    lit["material"]["data"].append({"type": "diffuse_light", "texture": len(lit["texture"]["data"]) - 1})
    lit["object"]["data"].append({"type": "sphere", "center": [6, 4, 6], "radius": 1.0, "material": len(lit["material"]["data"]) - 1})
    rule = lambda b, f: animation.moving_light_frame(b, f, 24, "{:03d}.png")
    (tmp_path / "lf").mkdir(); (tmp_path / "li").mkdir()
    animation.render_animation(rt, lit, rule, [0, 3, 6], str(tmp_path / "lf"), width=64, height=36, spp=2)
    r = animation.render_animation(rt, lit, rule, [0, 3, 6], str(tmp_path / "li"), width=64, height=36, spp=2, incremental=True)
    assert r["incremental_updates"] == 2
    for f in (0, 3, 6):
        a, b = np.asarray(Image.open(tmp_path / "lf" / f"{f:03d}.png")), np.asarray(Image.open(tmp_path / "li" / f"{f:03d}.png"))
        assert np.array_equal(a, b)
    assert not np.array_equal(np.asarray(Image.open(tmp_path / "lf" / "000.png")), np.asarray(Image.open(tmp_path / "lf" / "006.png")))
    bad = capi.ObjectDesc(); bad.type = 0; bad.material = 99; bad.xform = -1
    with pytest.raises(RtError, match="missing material"):
        rt.scene_update([bad], 0)
    rt.render(seed=1); rt.download()            # the context and its scene are still usable


def test_million_triangle_mesh_tile_lists_equal_traversal_at_4k(rt):
    """1 M triangles at 3840 x 2160, 128 spp (1 G paths): bounce 0 resolved from the tile candidate lists (64 candidates per 8x4 tile)
    gives the same image and ray count as bounce 0 through the tree — the check that exposed the grazing-triangle acceptance bug."""
    from rt_b200 import scenes
    ts = scenes.mesh_scene(708, 708, 3840, 2160, 128)
    rt.upload(ts)
    rt.stats_reset(); rt.render(seed=0x2023); a, _, sa = rt.download()
    rt.stats_reset(); rt.render(seed=0x2023, flags=capi.RT_RENDER_NO_TILE_LISTS); b, _, sb = rt.download()
    assert sa["rays"] == sb["rays"]
    assert np.array_equal(a, b), int((a != b).any(-1).sum())
