#!/usr/bin/env python3
"""bench.py — headline measurement of the render path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]
  (N > 1: launched by torchrun, one rank per GPU; strong scaling: the spp of one image are split across ranks and
   the float accumulation buffers are summed to rank 0 with one NCCL reduce over NVLink.)

One "step" = one complete render of the workload image (all samples per pixel) with the scene already resident in
HBM; `e2e` = the same step through the reference-facing call sequence with HOST buffers (JSON text -> upload + LBVH
build -> render -> reduce -> download of float sums and RGB8).  Prints ONE JSON line (contract in the task statement).
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

WORKLOADS = {
    # name: (scene file, width, height, spp) — BASELINE.json north_star target + configs
    "rtiow_1080p_1024spp": ("rtiow_486.json", 1920, 1080, 1024),
    "rtiow_c1_1200x800_10spp": ("rtiow_486.json", 1200, 800, 10),
    "basic_scene_1080p_1024spp": ("basic_scene.json", 1920, 1080, 1024),
    "dna000_1080p_1024spp": ("dna_000.json", 1920, 1080, 1024),
    "blue2_1080p_256spp": ("blue2.json", 1920, 1080, 256),
    "sample_scene_1600x900_512spp": ("sample_scene.json", 1600, 900, 512),
    # config 3 as BASELINE.json words it (boxes, rotate_y, image + checker textures, three emitters, OBJ mesh): our schema extensions
    "sample_scene_ext_1600x900_512spp": ("sample_scene_ext.json", 1600, 900, 512),
    # config 3 at its full sample count (4096 spp, 512 per GPU on 8)
    "sample_scene_1600x900_4096spp": ("sample_scene.json", 1600, 900, 4096),
    "sample_scene_ext_1600x900_4096spp": ("sample_scene_ext.json", 1600, 900, 4096),
    # config 4: procedural ~1M-triangle mesh (scenes.mesh_scene, seeded) under the GPU LBVH; "@" = tables, not a JSON file
    "mesh1m_4k_1024spp": ("@mesh:708", 3840, 2160, 1024),
    "mesh1m_1080p_64spp": ("@mesh:708", 1920, 1080, 64),
    # config 5: blue2.py's moving-cylinder animation, frames dealt round-robin to the ranks (no collective)
    "blue2_anim_240f_1080p_256spp": ("@anim:blue2:240", 1920, 1080, 256),
    "blue2_anim_16f_1080p_64spp": ("@anim:blue2:16", 1920, 1080, 64),
}
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "cmake_cpu_mt")
REF_GPU_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_harness")


def flop_model(st):
    """Algorithmic FLOPs of SURVEY.md §8(d); extend = box + primitive tests, shade = shade + generate."""
    ext = (24 * st["n_box_tests"] + 23 * st["n_sphere_reject"] + 47 * st["n_sphere_accept"] + 6 * st["n_rect_tests"] +
           62 * st["n_cyl_reject"] + 130 * st["n_cyl_accept"] + 48 * st["n_tri_tests"] + 24 * st["n_boxprim_tests"])
    shade = 60 * st["n_shade"] + 30 * st["n_generate"]
    return ext, shade


class ClockSampler:
    """SM clock, power and throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe: the nvidia-smi clocks
    line).  Read through NVML in this process every 50 ms — the same counters nvidia-smi prints, without starting a process
    inside the timed region (its start-up takes driver locks for tens of ms and showed up as 45 -> 60 ms steps on the
    launch-bound workloads); `nvidia-smi -lms` is the fallback when the NVML binding is missing."""

    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        self.rows = []          # (sm_mhz, sm_max_mhz, power_w, reasons bitmask)
        self.proc = None
        self.index = index
        self.source = None
        self._stop = threading.Event()
        self._thread = None

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def start(self):
        try:
            nv, h = self._nvml_handle()
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)

            def loop():
                while not self._stop.is_set():
                    try:
                        try:
                            rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                        except Exception:
                            rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append((float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), float(mx),
                                          nv.nvmlDeviceGetPowerUsage(h) / 1000.0, int(rs)))
                    except Exception:
                        pass
                    self._stop.wait(0.05)
            self.source = "nvml"
            self._thread = threading.Thread(target=loop, daemon=True)
            self._thread.start()
            return
        except Exception:
            self.source = None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            c = [x.strip() for x in line.split(",")]
            try:
                mask = sum(bit for k, (_, bit) in enumerate(self.REASONS) if len(c) > 3 + k and c[3 + k].lower().startswith("active"))
                self.rows.append((float(c[0]), float(c[1]), float(c[2]), mask))
            except (ValueError, IndexError):
                pass

    def stop(self):
        if self.source is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        if self.source == "nvml":
            time.sleep(0.06)
            self._stop.set()
            self._thread.join(timeout=1.0)
        else:
            time.sleep(0.25)
            self.proc.terminate()
        sm = sorted(r[0] for r in self.rows)
        reasons = [n for n, bit in self.REASONS if any(r[3] & bit for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max((r[1] for r in self.rows), default=None), "reasons": reasons,
                "power_w_max": max((r[2] for r in self.rows), default=None), "samples": len(self.rows), "source": self.source}


def scene_text(name, width, height, spp):
    d = json.load(open(os.path.join(GOLDEN, name)))
    if "aspect_ratio" not in d["camera"]:
        d["camera"]["aspect_ratio"] = d["width"] / d["height"]     # keep the scene's own aspect when the size is overridden
    d["width"], d["height"], d["samples_per_pixel"] = width, height, spp
    return json.dumps(d)


def make_upload(rt, workload):
    """Returns (upload(renderer), description dict): JSON text through rt_scene_upload_json_text, or flat tables through
    rt_scene_upload for the procedural mesh workloads."""
    scene, W, H, spp = WORKLOADS[workload]
    if scene.startswith("@mesh"):
        from rt_b200 import scenes
        n = int(scene.split(":")[1])
        ts = scenes.mesh_scene(n, n, W, H, spp)
        desc = {"max_depth": ts.desc.max_depth, "objects": int(ts.desc.n_objects), "h2d": int(ts.objects.nbytes) + 32 * 8, "tables": ts,
                "gradient": ts.desc.background_mode == 1}
        return (lambda r: r.upload(ts)), desc
    text = scene_text(scene, W, H, spp)
    d = json.loads(text)
    h2d = 64 * len(d["object"]["data"]) + 32 * (len(d["material"]["data"]) + len(d["texture"]["data"])) + len(text)
    return (lambda r: r.upload_json_text(text, GOLDEN)), {"max_depth": d["max_depth"], "objects": len(d["object"]["data"]), "h2d": h2d, "text": text,
                                                          "gradient": d.get("background_mode") == "sky_gradient"}


def run_reference_binary(width, height, spp, procs):
    """Runs `procs` concurrent copies of the UNMODIFIED reference cmake-cpu-version binary (it is single-threaded:
    cmake-cpu-version/main.cpp has no omp pragma).  Returns wall seconds for all of them."""
    dirs = [tempfile.mkdtemp(prefix="refcpu_") for _ in range(procs)]
    t0 = time.time()
    ps = [subprocess.Popen([REF_BIN, "-w", str(width), "-h", str(height), "-spp", str(spp), "-d", "50"], cwd=d,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for d in dirs]
    rcs = [p.wait() for p in ps]
    dt = time.time() - t0
    for d in dirs:
        shutil.rmtree(d, ignore_errors=True)
    if any(rcs):
        raise RuntimeError(f"reference binary failed: {rcs}")
    return dt


def _still_scene(workload):
    """(scene file or @mesh tag, W, H, spp); an animation workload is represented by its base scene (frame rule not applied)."""
    scene, W, H, spp = WORKLOADS[workload]
    if scene.startswith("@anim"):
        scene = {"blue2": "blue2.json", "blue": "blue.json", "dna": "dna_000.json"}[scene.split(":")[1]]
    return scene, W, H, spp


def oracle_sample(workload, threads, w, h, s):
    """The CPU restatement (oracle/, brute-force closest hit like the reference's hittable_list) on a bounded sample."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    scene, W, H, spp = _still_scene(workload)
    if scene.startswith("@mesh"):
        rt = ge.load_package()
        from rt_b200 import scenes
        n = int(scene.split(":")[1])
        sc = scenes.mesh_scene(n, n, w, h, s)
        sc.desc.camera.aspect = W / H
    else:
        sc = orc.load_scene(json.loads(scene_text(scene, W, H, spp)), width=w, height=h, spp=s)
    t0 = time.time(); orc.render(sc, threads=threads); dt = time.time() - t0
    return w * h * s, dt, f"oracle port {w}x{h}x{s}spp, {threads} thread(s), brute-force closest hit", "port"


def reference_sample(workload, procs):
    """A bounded sample of the workload for the CPU arm: same scene and camera, reduced size (the reference keeps the
    16:9 camera regardless of -w/-h, main.cpp:64,94).  Returns (samples, seconds, description, kind)."""
    scene, W, H, spp = _still_scene(workload)
    if scene.startswith("@mesh"):      # the reference has no C++ mesh path at all: the oracle port is the CPU statement
        return oracle_sample(workload, procs, 32, 18, 1)
    if scene == "rtiow_486.json" and os.path.exists(REF_BIN):
        w, h, s = 480, 270, 8          # 8 spp: the reference's per-pixel by-value copy of the world (main.cpp:45-55) is amortised
        dt = run_reference_binary(w, h, s, procs)
        return procs * w * h * s, dt, f"{procs} x cmake_cpu_mt -w {w} -h {h} -spp {s} -d 50 (unmodified reference binary, 1 thread each)", "reference"
    if not os.path.exists(REF_GPU_HARNESS):
        return oracle_sample(workload, procs, 480, 270, 4)
    # JSON scenes: the reference gpu-version headers compiled for the host (oracle/_ref/ref_gpu_harness), `procs` threads
    w, h, s = 480, 270, 16
    d = json.loads(scene_text(scene, w, h, s))
    d.pop("aspect_ratio", None); d["camera"].pop("aspect_ratio", None)
    with tempfile.TemporaryDirectory() as td:
        fs = os.path.join(td, "s.json"); json.dump(d, open(fs, "w"))
        t0 = time.time()
        subprocess.run([REF_GPU_HARNESS, "render", fs, "1", os.path.join(td, "o.bin"), str(procs)], check=True, capture_output=True)
        dt = time.time() - t0
    return w * h * s, dt, f"ref_gpu_harness render {w}x{h}x{s}spp, {procs} threads (reference gpu-version headers on the host)", "reference"


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    procs = os.cpu_count() or 1
    scene, W, H, spp = WORKLOADS[args.workload]
    for _ in range(args.warmup):
        reference_sample(args.workload, procs)
    tot_s, tot_t, desc, kind = 0, 0.0, "", "port"
    for _ in range(args.steps):
        n, dt, desc, kind = reference_sample(args.workload, procs)
        tot_s += n; tot_t += dt
    v = tot_s / tot_t
    print(json.dumps({
        "impl": "reference", "metric": "samples_per_second", "value": v, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64" if scene == "rtiow_486.json" else "f32", "data": "synthetic",
        "config": {"workload": args.workload, "scene": scene, "width": W, "height": H, "spp": spp, "note": "CPU arm renders a bounded sample of this workload per step"},
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": procs, "kind": kind, "sample": desc},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


def main_animation(args):
    """Config 5.  One step = every frame of the animation once; rank r renders frames f with f % world == r, each frame
    through the full host path (frame rule -> JSON text -> upload + LBVH build -> render -> RGB8 download -> PNG file).
    `value` = samples / device time (CUDA events around each frame's render), `e2e` = samples / wall time of the step."""
    import tempfile as tf
    import torch
    import torch.distributed as dist
    rt = ge.load_package()
    from rt_b200 import animation, multi_gpu
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    scene, W, H, spp = WORKLOADS[args.workload]
    _, name, n_frames = scene.split(":")
    base_file, fn, _ = animation.FRAME_RULES[name]
    base = json.load(open(os.path.join(GOLDEN, base_file)))
    frames = multi_gpu.frames_for_rank(int(n_frames), rank, world)
    r = rt.Renderer(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        with tf.TemporaryDirectory() as td:
            return animation.render_animation(r, base, fn, frames, td, W, H, spp, base_dir=GOLDEN, skip_existing=False)
    for _ in range(args.warmup):
        animation.render_animation(r, base, fn, frames[:2], None, W, H, spp, base_dir=GOLDEN)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    dev_ms, rays, launches = 0.0, 0, 0
    for _ in range(args.steps):
        res = step()
        dev_ms += res["device_ms"]; rays += res["rays"]; launches += res["launches"]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([dev_ms, wall * 1e3, float(rays)], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, wall_ms, rays = float(mx[0]), float(mx[1]), float(sm[2])
    else:
        wall_ms = wall * 1e3
    if rank == 0:
        total = int(n_frames) * W * H * spp
        line = {"metric": "samples_per_second", "value": total * args.steps / (dev_ms * 1e-3), "unit": "samples/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": args.workload, "scene": base_file, "frames": int(n_frames), "width": W, "height": H, "spp": spp,
                           "partition": f"frames dealt round-robin over {world} GPU(s), no collective",
                           "l2": "wavefront queues (GBs) exceed the 126 MB L2 every pass; no explicit flush needed"},
                "mrays_per_s": rays / (dev_ms * 1e-3) / 1e6,
                "e2e": {"value": total * args.steps / (wall_ms * 1e-3), "unit": "samples/s", "ms_per_step": wall_ms / args.steps,
                        "h2d_bytes_per_step": int(n_frames) * 2048, "d2h_bytes_per_step": int(n_frames) * W * H * 3,
                        "note": "frame rule + JSON text + upload/LBVH build + render + RGB8 download + PNG encode/write (8 host threads)"},
                "gpu_launches": int(launches / args.steps), "clocks": clocks, "roofline": None}
        print(json.dumps(line))
    r.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="rtiow_1080p_1024spp", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sort", action="store_true", help="sort shade chunks by material class (RT_RENDER_SORT)")
    ap.add_argument("--max-paths", type=int, default=0)
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)
    if WORKLOADS[args.workload][0].startswith("@anim"):
        return main_animation(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    rt = ge.load_package()
    from rt_b200 import capi, multi_gpu

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    scene, W, H, spp = WORKLOADS[args.workload]
    upload, sdesc = make_upload(rt, args.workload)
    stream = torch.cuda.Stream()
    s0, s1 = multi_gpu.spp_slice(spp, rank, world)        # this rank's whole sample chunks (rt_rank_chunks); the library deals them itself
    flags = capi.RT_RENDER_SORT if args.sort else 0       # timed region: the default schedule (two passes in flight)

    with torch.cuda.stream(stream):
        r = rt.Renderer(local, stream=stream.cuda_stream)
        multi_gpu.init_comm(r)        # the LIBRARY's NCCL communicator (csrc/rt_multi.cu); torch.distributed only carries the unique id
        upload(r)

        def step_resident():
            # every rank renders its chunks into per-chunk buffers; the library gathers them on rank 0 over NCCL and adds them in
            # chunk order (bit-identical image for any rank count)
            r.render_shard(seed=0x2023, spp_count=spp, flags=flags, max_paths_per_pass=args.max_paths)
            r.gather_image(0)

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        fp32_peak = r.probe_fp32_peak() if rank == 0 else 0.0
        # event counts for the flop model (deterministic: counter-based RNG); outside the timed region
        counts = None
        if rank == 0:
            r.stats_reset()
            r.render(seed=0x2023, spp_begin=0, spp_count=max(1, spp // 16), flags=capi.RT_RENDER_COUNT_EVENTS)
            _, _, counts = r.download(want_sum=False, want_rgb8=False)
            scale = spp / max(1, spp // 16) / world      # counts scale linearly with the number of samples
        for _ in range(args.warmup):
            step_resident()
        barrier()
        r.stats_reset()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            step_resident()
        e1.record(stream)
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        ms = torch.tensor([e0.elapsed_time(e1)], device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_step = float(ms.item()) / args.steps
        _, _, st = r.download(want_sum=False, want_rgb8=False)       # rays and launches of the timed region
        # per-stage times: one extra step right after the timed region with a CUDA-event pair around every launch, which
        # needs one pass at a time (RT_RENDER_TIME_STAGES serialises the two pass streams)
        r.stats_reset()
        r.render(seed=0x2023, spp_begin=s0, spp_count=s1 - s0, total_spp=spp, flags=flags | capi.RT_RENDER_TIME_STAGES | capi.RT_RENDER_KEEP_CHUNKS,
                 max_paths_per_pass=args.max_paths)
        _, _, sg = r.download(want_sum=False, want_rgb8=False)
        # the same with the queue schedule kept to the last bounce (no tail stage): every ray passes through k_extend / k_shade, which is
        # what the event counters of the counted run describe — the roofline fractions are taken on THIS step
        r.stats_reset()
        r.render(seed=0x2023, spp_begin=s0, spp_count=s1 - s0, total_spp=spp, max_paths_per_pass=args.max_paths,
                 flags=flags | capi.RT_RENDER_TIME_STAGES | capi.RT_RENDER_KEEP_CHUNKS | capi.RT_RENDER_NO_TAIL)
        _, _, sq = r.download(want_sum=False, want_rgb8=False)
        barrier()

        # ---- e2e: host JSON text -> upload (+BVH build) -> render -> reduce -> download to host, every step ----
        pin_sum = r.pinned_empty((H, W, 3), np.float32) if rank == 0 else None      # page-locked download targets
        pin_rgb8 = r.pinned_empty((H, W, 3), np.uint8) if rank == 0 else None

        def step_e2e():
            upload(r)
            r.render_shard(seed=0x2023, spp_count=spp, flags=0, max_paths_per_pass=args.max_paths)
            r.gather_image(0)
            if rank == 0:
                return r.download(gamma=0, out_sum=pin_sum, out_rgb8=pin_rgb8)
            r.synchronize()
            return None
        step_e2e()
        barrier()
        t0 = time.perf_counter()
        e0.record(stream)
        n_e2e = max(2, args.steps)            # as many end-to-end steps as timed steps: one host hiccup (page faults, a busy core) is averaged over them
        for _ in range(n_e2e):
            out = step_e2e()
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
        ms2 = torch.tensor([max(e0.elapsed_time(e1), wall * 1e3)], device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
        ms_e2e = float(ms2.item()) / n_e2e

    if rank == 0:
        # integrity: the render is bit-reproducible (counter-based RNG, fixed-order accumulation, chunk-ordered multi-GPU sum), so the
        # float image of the last end-to-end step must hash to the committed value of this workload — on ANY number of GPUs
        import hashlib
        image_sha = hashlib.sha256(np.ascontiguousarray(out[0]).tobytes()).hexdigest()
        hpath = os.path.join(GOLDEN, "image_hashes.json")
        known = json.load(open(hpath)) if os.path.exists(hpath) else {}
        want_sha = known.get(args.workload)
        image_check = "no committed hash" if want_sha is None else ("ok" if want_sha == image_sha else "MISMATCH")
        total_samples = W * H * spp
        value = total_samples / (ms_step * 1e-3)
        rays = st["rays"] / args.steps * world            # rank 0 traced 1/world of them
        ext_flops, shade_flops = flop_model(counts)
        ext_flops *= scale; shade_flops *= scale
        ms_extend = sq["ms_extend"]
        ms_other = sq["ms_generate"] + sq["ms_shade"] + sq["ms_accumulate"]
        ms_tail = sg.get("ms_tail", 0.0)
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        ach = ext_flops / (ms_extend * 1e-3) / 1e12 if ms_extend > 0 else 0.0
        rays_rank, samples_rank = st["rays"] / args.steps, total_samples / world
        h2d = sdesc["h2d"]
        # DRAM traffic of the dominant kernel per launch, from the committed ncu capture (bytes per ray x rays per launch)
        traffic, traffic_note = None, ""
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        depth = int(sdesc["max_depth"])
        bounce0_in_generate = counts.get("n_generate_leaf_tests", 0) > 0       # primary hits resolved from the tile candidate lists
        n_ext_launches = max(1, int(sq.get("launches_extend", 1)))
        # rays of the extend launches: the counted run (no tail stage there) saw every ray; the timed schedule hands the paths alive at
        # the tail bounce to k_tail, whose rays and time are reported separately (stage "tail") and are NOT in this roofline
        ext_rays = rays_rank - (samples_rank if bounce0_in_generate else 0.0)
        # algorithmic bytes of the queue stages on this rank: generate W32/sample (ray_o, ray_d; camera paths carry
        # no throughput record) + W8/sample when it also resolves the bounce-0 hits; shade R56/ray minus the 16-byte throughput
        # record at bounce 0, + W48 per survivor (rays - samples) + W16 per finished path; accumulate R16/sample (+12 B/pixel RMW
        # per pass, ignored).  Paths of empty camera tiles move fewer bytes than this (not subtracted: no such tile in RTIOW).
        # Camera paths of empty tiles (their beam touches no leaf box) never enter the queues: under a constant background they move no
        # bytes at all, under the sky gradient only their 16-byte radiance record (written by generate, read by accumulate).
        skipped = float(sq.get("paths_skipped", 0))
        gradient = sdesc.get("gradient", False)
        queued = samples_rank - skipped
        queue_bytes = ((32.0 + (8.0 if bounce0_in_generate else 0.0)) * queued + 56.0 * (rays_rank - skipped) - 16.0 * queued
                       + 48.0 * (rays_rank - samples_rank) + 16.0 * queued + 16.0 * queued + (32.0 * skipped if gradient else 0.0))
        if sq["ms_shade"] == 0 and sq["ms_extend"] == 0:      # no queue stage ran (empty scene, or the one-kernel path of all-emitter scenes):
            queue_bytes = 2.0 * 12.0 * W * H * 8              # only the chunk buffers are written and folded
        hbm_ach = queue_bytes / (ms_other * 1e-3) / 1e9 if ms_other > 0 else 0.0
        if args.workload == "rtiow_1080p_1024spp" and os.path.exists(tpath):
            tj = json.load(open(tpath))
            traffic = tj["k_extend_dram_bytes_per_ray"] * ext_rays / n_ext_launches
            traffic_note = "; traffic = %.1f B/ray (ncu, profiles/traffic.json) x mean rays per extend launch, bytes" % tj["k_extend_dram_bytes_per_ray"]
        # the tail stage (closest hit + shade of every bounce from the hand-over bounce on, one launch per pass): its rays are the ones
        # neither generate nor an extend launch of the default schedule traced; numerator = their share of the counted run's
        # algorithmic FLOPs (a per-ray average over all bounces: late bounces walk longer, so this understates the tail slightly)
        tail_rays = max(0.0, rays_rank - (samples_rank if bounce0_in_generate else 0.0) - float(sg.get("rays_extend", 0)))
        tail_flops = ext_flops * (tail_rays / ext_rays if ext_rays > 0 else 0.0) + shade_flops * (tail_rays / rays_rank if rays_rank > 0 else 0.0)
        tail_ach = tail_flops / (ms_tail * 1e-3) / 1e12 if ms_tail > 0 else 0.0
        stages = {"generate": sg["ms_generate"], "extend": sg["ms_extend"], "shade": sg["ms_shade"], "tail": ms_tail, "accumulate": sg["ms_accumulate"]}
        line = {
            "metric": "samples_per_second", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "scene": scene, "width": W, "height": H, "spp": spp, "max_depth": sdesc["max_depth"],
                       "objects": sdesc["objects"], "partition": f"8 sample chunks dealt to {world} GPU(s); chunk sums gathered on rank 0 by the library (NCCL send/recv) and added in chunk order",
                       "l2": "wavefront queues (GBs) exceed the 126 MB L2 every pass; no explicit flush needed"},
            "mrays_per_s": rays / (ms_step * 1e-3) / 1e6, "rays_per_sample": rays / total_samples,
            "stage_ms_per_step": {"generate": sg["ms_generate"], "extend": sg["ms_extend"], "shade": sg["ms_shade"], "accumulate": sg["ms_accumulate"],
                                  "tail": ms_tail, "serial_total": sg["ms_total"],
                                  "note": "one extra step with per-launch CUDA events, one pass at a time; the timed steps keep two passes in flight",
                                  "queue_schedule": {"generate": sq["ms_generate"], "extend": sq["ms_extend"], "shade": sq["ms_shade"],
                                                     "accumulate": sq["ms_accumulate"], "serial_total": sq["ms_total"],
                                                     "note": "RT_RENDER_NO_TAIL: every bounce through the queues; the rooflines are taken on this step"}},
            "roofline": {"bound": "fp32", "kernel": "k_extend", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": ach / fp32_peak if fp32_peak else None, "traffic": traffic,
                         "note": "algorithmic FLOPs (SURVEY §8(d) model, device event counters of the extend stage only) / CUDA-event time of the extend launches "
                                 "of one queue-schedule step (RT_RENDER_NO_TAIL) run right after the timed region; peak = FP32 FMA-chain microbenchmark measured in this run (of measured)" + traffic_note},
            "roofline_tail": {"bound": "fp32", "kernel": "k_tail", "achieved": tail_ach, "peak": fp32_peak, "unit": "TFLOP/s",
                              "frac": tail_ach / fp32_peak if fp32_peak else None, "rays_share": tail_rays / rays_rank if rays_rank > 0 else 0.0,
                              "note": "closest hit + shade of every bounce from the hand-over bounce on, one launch per pass: the share of the counted run's extend + shade "
                                      "FLOPs that belongs to the rays this stage traced / its CUDA-event time in the default schedule"},
            "largest_stage": max(stages, key=stages.get),
            "roofline_hbm": {"bound": "hbm", "kernel": "generate+shade+accumulate", "achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s",
                             "frac": hbm_ach / hbm_peak, "note": ("104 B/ray (+ 8 B/sample when generate writes the bounce-0 hit records; DESIGN.md §6) / CUDA-event time of those stages; peak from MEASURED_PEAKS.json (of measured)" if peaks else "of fallback")},
            "e2e": {"value": total_samples / (ms_e2e * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": W * H * 3 * 4 + W * H * 3, "ms_per_step": ms_e2e},
            "gpu_launches": int(st["launches"] / args.steps), "clocks": clocks,
            "image_sha256": image_sha, "image_check": image_check,
        }
        if world == 1 and not args.no_cpu_baseline:
            n, dt, desc, kind = reference_sample(args.workload, 1)
            line["cpu_baseline"] = {"value": n / dt, "unit": "samples/s", "cores": 1, "kind": kind, "sample": desc.replace("1 x ", "")}
        print(json.dumps(line))
    r.close()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and image_check == "MISMATCH":
        sys.stderr.write(f"bench.py: image hash {image_sha} differs from the committed {want_sha} for {args.workload}\n")
        sys.exit(3)


if __name__ == "__main__":
    main()
