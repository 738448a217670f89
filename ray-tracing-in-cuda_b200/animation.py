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


def frames_for_rank(n_frames, rank, world):
    """Frame partition of blue.py:23-32: frame f renders on GPU f % 8."""
    return [f for f in range(n_frames) if f % world == rank]


# ---- in-process animation driver (replaces "write 360 JSON files, start one process per frame per GPU") ----
FRAME_RULES = {
    # name: (base scene file, frame function, number of frames the reference script renders)
    "dna": ("basic_scene.json", dna_frame, 360),         # dna.py:17 (it writes 360 scenes; its render loop covers them all)
    "blue": ("blue.json", blue_frame, 360),              # blue.py:14
    "blue2": ("blue2.json", blue2_frame, 180),           # blue2.py:14
}


def render_animation(renderer, base_scene, frame_fn, frames, out_dir=None, width=0, height=0, spp=0, seed=0x2023,
                     writer_threads=8, skip_existing=True, base_dir=None):
    """Renders `frames` (an iterable of frame numbers — this rank's share, see frames_for_rank) on one Renderer.

    Per frame: scene dict from the reference's rule -> JSON text -> rt_scene_upload_json_text (parse, flatten, LBVH build
    on the device, ~0.2 ms) -> rt_render -> rt_download (tonemapped RGB8, rows flipped on the device) -> PNG encoding on a
    pool of host threads, overlapped with the next frame's rendering.  Frames whose PNG already exists are skipped
    (frame-level resume; the reference restarts from frame 0).  Returns a dict with counts and timings."""
    import json
    import os
    import time
    from concurrent.futures import ThreadPoolExecutor

    from .renderer import write_png

    t0 = time.perf_counter()
    done, skipped, samples, rays, launches, device_ms, pending = 0, 0, 0, 0, 0, 0.0, []
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
        info = renderer.upload_json_text(json.dumps(scene), base_dir)
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
    return {"frames": done, "skipped": skipped, "samples": samples, "rays": rays, "launches": launches, "seconds": dt, "device_ms": device_ms}
