"""Per-frame scene rules of the reference's animation drivers, restated as pure functions.

gpu-version/dna.py:17-85 (three DNA helices of emissive spheres + cylinders filled into basic_scene.json,
angle = frame degrees), gpu-version/blue.py:16-22 (+1 deg per frame on every cylinder of blue.json, 360
frames) and gpu-version/blue2.py:16-22 (+2 deg per frame, 180 frames).  The reference writes one JSON per
frame and shells out to ./build/parallel_compute once per frame per GPU; here the frames are generated in
memory and handed to rt_scene_upload_json_text, partitioned frame -> GPU (frame % n_gpus) by the caller.
"""
import copy
import math


def dna_frame(basic_scene, frame, output_pattern="./build/output/output_{:03d}.png"):
    angle = frame * 1                      # dna.py:17: range(0, 360, 60 // framerate) with framerate = 60
    scene = copy.deepcopy(basic_scene)
    scene["output_file"] = output_pattern.format(frame)
    tex, mat, obj = [], [], []
    num_object, space = 5, 5
    for i, _id in enumerate(range(-num_object * 3, num_object * 3)):       # dna.py:29-54
        tex += [{"type": "solid_color", "color": [232 / 256, 209 / 256, 209 / 256]},
                {"type": "solid_color", "color": [232 / 256, 209 / 256, 209 / 256]},
                {"type": "solid_color", "color": [202 / 256, 202 / 256, 224 / 256]}]
        mat += [{"type": "diffuse_light", "texture": i * 3 + k} for k in range(3)]
    for offset in range(3):                                                    # dna.py:55-84
        for i, _id in enumerate(range(-num_object, num_object)):
            theta = (36 * (_id + num_object) + angle) / 180 * math.pi
            xoffset = offset * space - space
            zoffset = math.fabs(offset - 1) * -20 + 20
            obj.append({"type": "sphere", "center": [2.5 * math.cos(theta) + xoffset, _id, 2.5 * math.sin(theta) + zoffset],
                        "radius": 0.5, "material": i * 3 + 0})
            obj.append({"type": "sphere", "center": [2.5 * math.cos(theta + math.pi) + xoffset, _id,
                                                     2.5 * math.sin(theta + math.pi) + zoffset],
                        "radius": 0.5, "material": i * 3 + 1})
            obj.append({"type": "cylinder", "radius": 0.3, "zmin": -2.18, "zmax": 2.18, "material": i * 3 + 2,
                        "translate": [0 + xoffset, _id, 0 + zoffset],
                        "rotate": {"axis": [0, 1, 0], "angle": 36 * -(_id + num_object) + 90 + angle}})
    scene["texture"] = {"data": tex}
    scene["material"] = {"data": mat}
    scene["object"] = {"data": obj}
    return scene


def rotate_cylinders_frame(base_scene, frame, step_deg, output_pattern):
    """blue.py (step 1, 360 frames) / blue2.py (step 2, 180 frames): frame f has every cylinder at base + step*(f+1)."""
    scene = copy.deepcopy(base_scene)
    for item in scene["object"]["data"]:
        if item["type"] == "cylinder":
            item["rotate"]["angle"] = item["rotate"]["angle"] + step_deg * (frame + 1)
    scene["output_file"] = output_pattern.format(frame)
    return scene


def blue_frame(blue_scene, frame):
    return rotate_cylinders_frame(blue_scene, frame, 1, "./build/output/blue/frame_{:03d}.png")


def blue2_frame(blue2_scene, frame):
    return rotate_cylinders_frame(blue2_scene, frame, 2, "./build/output/blue2/{:03d}.png")


def moving_light_frame(base_scene, frame, n_frames=120, output_pattern="./build/output/light/{:03d}.png"):
    """The reference's open to-do item "场景不变，移动光源" (scene fixed, light moving; README.md:14-17): every emissive sphere circles the
    vertical axis through the camera's look-at point, one turn over n_frames; nothing else changes."""
    scene = copy.deepcopy(base_scene)
    cx, _, cz = scene["camera"]["lookat"]
    a = 2.0 * math.pi * frame / n_frames
    ca, sa = math.cos(a), math.sin(a)
    mats = scene["material"]["data"]
    for item in scene["object"]["data"]:
        if item["type"] == "sphere" and mats[item["material"]]["type"] == "diffuse_light":
            x, y, z = item["center"]
            dx, dz = x - cx, z - cz
            item["center"] = [cx + ca * dx + sa * dz, y, cz - sa * dx + ca * dz]
    scene["output_file"] = output_pattern.format(frame)
    return scene


def frames_for_rank(n_frames, rank, world):
    """Frame partition of blue.py:23-32: frame f renders on GPU f % 8."""
    return [f for f in range(n_frames) if f % world == rank]


# ---- in-process animation driver (replaces "write 360 JSON files, start one process per frame per GPU") ----
FRAME_RULES = {
    # name: (base scene file, frame function, number of frames the reference script renders)
    "dna": ("basic_scene.json", dna_frame, 360),         # dna.py:17 writes 360 scenes (its own render loop then takes range(3) of them)
    "blue": ("blue.json", blue_frame, 360),              # blue.py:14
    "blue2": ("blue2.json", blue2_frame, 180),           # blue2.py:14
    "light": ("sample_scene_ext.json", moving_light_frame, 120),     # README.md:14-17 (no script in the reference: our rule)
}


# ---- frame-to-frame updates without scene text (rt_scene_update) ----
_INSTANCE_TYPES = ("sphere", "xy_rect", "xz_rect", "yz_rect", "box")
_OBJ_TYPE = {"sphere": 0, "xy_rect": 1, "xz_rect": 2, "yz_rect": 3, "cylinder": 4, "box": 5}


def _has_xform(o):
    return o["type"] == "cylinder" or (o["type"] in _INSTANCE_TYPES and ("rotate" in o or "translate" in o))


def _object_desc(o, xform_index):
    """capi.ObjectDesc of one JSON object (field order of include/rt.h; the JSON reader's own conversion, csrc/rt_json.cpp)."""
    from . import capi
    q = capi.ObjectDesc()
    q.type, q.material, q.xform = _OBJ_TYPE[o["type"]], int(o["material"]), xform_index
    t = o["type"]
    if t == "sphere":
        vals = list(o["center"]) + [o["radius"]]
    elif t.endswith("_rect"):
        a, b = t[0], t[1]
        vals = [o[a + "0"], o[a + "1"], o[b + "0"], o[b + "1"], o["k"]]
    elif t == "cylinder":
        vals = [o["radius"], o["zmin"], o["zmax"]]
    else:
        vals = list(o["min"]) + list(o["max"])
    for k, v in enumerate(vals):
        q.p[k] = v
    return q


def scene_delta(prev, cur):
    """What rt_scene_update needs to turn uploaded scene `prev` into `cur`: (first_object, [object dicts], first_xform, [objects whose
    transform changed...]) as index ranges — or None when anything but objects of the simple types moved (then: full upload).
    Object table index = JSON index and transform index = running count of transformed objects, which holds when the scene has
    no triangle / mesh objects (those expand or bake their transforms in the reader)."""
    strip = lambda d: {k: v for k, v in d.items() if k not in ("object", "output_file")}
    po, co = prev["object"]["data"], cur["object"]["data"]
    if strip(prev) != strip(cur) or len(po) != len(co):
        return None
    if any(o["type"] not in _OBJ_TYPE for o in co) or any(a["type"] != b["type"] or _has_xform(a) != _has_xform(b) for a, b in zip(po, co)):
        return None
    changed = [i for i, (a, b) in enumerate(zip(po, co)) if a != b]
    if not changed:
        return (0, 0, 0, 0)
    xi, k = {}, 0
    for i, o in enumerate(co):
        if _has_xform(o):
            xi[i] = k; k += 1
    o0, o1 = changed[0], changed[-1] + 1
    xs = [xi[i] for i in range(o0, o1) if i in xi]
    return (o0, o1, xs[0] if xs else 0, xs[-1] + 1 if xs else 0)


def apply_delta(renderer, cur, delta):
    o0, o1, x0, x1 = delta
    if o1 == o0:
        return
    objs = cur["object"]["data"]
    xi, k = {}, 0
    for i, o in enumerate(objs):
        if _has_xform(o):
            xi[i] = k; k += 1
    inv = {v: i for i, v in xi.items()}
    descs = [_object_desc(objs[i], xi.get(i, -1)) for i in range(o0, o1)]
    xforms = [renderer.make_xform(objs[inv[j]].get("rotate"), objs[inv[j]].get("translate")) for j in range(x0, x1)]
    renderer.scene_update(descs, o0, xforms, x0)


def render_animation(renderer, base_scene, frame_fn, frames, out_dir=None, width=0, height=0, spp=0, seed=0x2023,
                     writer_threads=8, skip_existing=True, base_dir=None, incremental=False):
    """Renders `frames` (an iterable of frame numbers — this rank's share, see frames_for_rank) on one Renderer.

    Per frame: scene dict from the reference's rule -> JSON text -> rt_scene_upload_json_text (parse, flatten, LBVH build
    on the device, ~0.2 ms) -> rt_render -> rt_download (tonemapped RGB8, rows flipped on the device) -> PNG encoding on a
    pool of host threads, overlapped with the next frame's rendering.  Frames whose PNG already exists are skipped
    (frame-level resume; the reference restarts from frame 0).  incremental: a frame that differs from the previous one only in
    objects of the simple types (moved / rotated spheres, rects, cylinders, boxes) is sent as table entries through rt_scene_update
    instead of JSON text — same pixels (the transforms come from rt_make_xform, the reader's own routine).
    Returns a dict with counts and timings."""
    import json
    import os
    import time
    from concurrent.futures import ThreadPoolExecutor

    from .renderer import write_png

    t0 = time.perf_counter()
    done, skipped, samples, rays, launches, device_ms, pending = 0, 0, 0, 0, 0, 0.0, []
    prev_scene, updates = None, 0
    pool = ThreadPoolExecutor(max_workers=max(1, writer_threads)) if out_dir else None
    for f in frames:
        scene = frame_fn(base_scene, f)
        out = os.path.join(out_dir, os.path.basename(scene.get("output_file", f"frame_{f:04d}.png"))) if out_dir else None
        if out and skip_existing and os.path.exists(out):
            skipped += 1
            continue
        if "aspect_ratio" not in scene["camera"]:
            scene["camera"]["aspect_ratio"] = scene["width"] / scene["height"]      # keep the framing under size overrides
        if width:
            scene["width"] = width
        if height:
            scene["height"] = height
        if spp:
            scene["samples_per_pixel"] = spp
        renderer.stats_reset()
        delta = scene_delta(prev_scene, scene) if (incremental and prev_scene is not None) else None
        if delta is not None:
            apply_delta(renderer, scene, delta); info = renderer.info(); updates += 1
        else:
            info = renderer.upload_json_text(json.dumps(scene), base_dir)
        prev_scene = scene
        renderer.render(seed=seed)
        _, rgb8, st = renderer.download(want_sum=False, gamma=0)
        samples += info.width * info.height * info.samples_per_pixel
        rays += st["rays"]; launches += st["launches"]; device_ms += st["ms_total"]
        if pool:
            pending.append(pool.submit(write_png, out, rgb8))
        done += 1
    for p in pending:
        p.result()
    if pool:
        pool.shutdown()
    dt = time.perf_counter() - t0
    return {"frames": done, "skipped": skipped, "samples": samples, "rays": rays, "launches": launches, "seconds": dt, "device_ms": device_ms,
            "incremental_updates": updates}


# ---- frame-sequence assembly (README.md:14-17: "video from the frames"; the reference has no code for it) ----
def write_apng(png_paths, out_path, fps=30, loops=0):
    """Assembles PNG frames of equal size (as written by rt_write_png: 8-bit RGB, one IDAT stream) into ONE animated PNG (APNG: acTL +
    per-frame fcTL / fdAT chunks), in the order given.  No external encoder: the frames' compressed streams are re-chunked as they are.
    Plays in browsers and ffmpeg; a plain PNG reader sees the first frame."""
    import struct
    import zlib

    def chunks(data):
        pos = 8
        while pos + 12 <= len(data):
            n = struct.unpack(">I", data[pos:pos + 4])[0]
            yield data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n]
            pos += 12 + n

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xffffffff)
    frames = []
    for p in png_paths:
        data = open(p, "rb").read()
        if data[:8] != b"\x89PNG\r\n\x1a\n":
            raise ValueError(f"{p} is not a PNG file")
        ihdr, idat = None, b""
        for tag, body in chunks(data):
            if tag == b"IHDR":
                ihdr = body
            elif tag == b"IDAT":
                idat += body
        frames.append((ihdr, idat))
    if not frames:
        raise ValueError("no frames")
    if any(f[0] != frames[0][0] for f in frames):
        raise ValueError("frames differ in size or pixel format")
    w, h = struct.unpack(">II", frames[0][0][:8])
    out = [b"\x89PNG\r\n\x1a\n", chunk(b"IHDR", frames[0][0]), chunk(b"acTL", struct.pack(">II", len(frames), loops))]
    seq = 0
    for k, (_, idat) in enumerate(frames):
        out.append(chunk(b"fcTL", struct.pack(">IIIIIHHBB", seq, w, h, 0, 0, 1, int(fps), 0, 0))); seq += 1
        if k == 0:
            out.append(chunk(b"IDAT", idat))
        else:
            out.append(chunk(b"fdAT", struct.pack(">I", seq) + idat)); seq += 1
    out.append(chunk(b"IEND", b""))
    with open(out_path, "wb") as f:
        f.write(b"".join(out))
    return len(frames)


def write_frame_list(png_paths, out_path, fps=30):
    """ffmpeg concat-demuxer list of the frames in order (`ffmpeg -f concat -safe 0 -i frames.txt -pix_fmt yuv420p out.mp4`): the
    hand-over to an external video encoder, which this image does not have."""
    import os
    with open(out_path, "w") as f:
        for p in png_paths:
            f.write(f"file '{os.path.abspath(p)}'\nduration {1.0 / fps:.6f}\n")
    return len(png_paths)
