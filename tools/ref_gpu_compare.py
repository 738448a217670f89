"""Same-box GPU comparison (last row of "CPU baseline timing"): the UNMODIFIED reference
gpu-version/main.cu compiled for sm_100 (oracle/_ref/parallel_compute_ref, recipe: oracle/Makefile target `refgpu`)
against this library on the same JSON scene files, on the same B200.

The reference binary prints progress lines on (unbuffered) stderr: its render time is the wall time between the arrival of
"Start rendering" and "Finish rendering" (main.cu:499-508: the render<<<>>> launch and the cudaDeviceSynchronize after it);
ours is the wall time of rt_render + rt_synchronize on a resident scene, and — like for like — the whole-process wall time
of our `parallel_compute` CLI on the same file next to the reference's.

Usage (GPU box):  python tools/ref_gpu_compare.py [--out gpurun_out/ref_gpu_compare.json]
Test infrastructure: only this tool and bench.py execute anything under oracle/_ref.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "parallel_compute_ref")
OUR_CLI = os.path.join(ROOT, "ray-tracing-in-cuda_b200", "parallel_compute")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def reference_schema(scene):
    """Strip what the reference's parser cannot read (parser.hpp:143-200 knows solid_color only): a checker texture becomes
    its `even` colour; our extension keys are ignored by the reference and dropped here for clarity."""
    s = json.loads(json.dumps(scene))
    texs = s["texture"]["data"]
    for t in texs:
        if t.get("type") == "checker":
            even = texs[t["even"]]
            t.clear(); t.update({"type": "solid_color", "color": even["color"]})
    for k in ("background_mode", "sampling"):
        s.pop(k, None)
    s["camera"].pop("aspect_ratio", None)
    return s


def run_wall(cmd, cwd, timeout):
    """Returns (whole-process wall s, exit code, stderr tail, seconds between the 'Start rendering' and 'Finish rendering' lines)."""
    t0 = time.perf_counter()
    p = subprocess.Popen(cmd, cwd=cwd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    t_start = t_finish = None
    tail = []
    for raw in p.stderr:                      # the reference writes these lines with fprintf(stderr, ...): unbuffered
        now = time.perf_counter()
        line = raw.decode(errors="replace")
        tail.append(line)
        if "Start rendering" in line:
            t_start = now
        elif "Finish rendering" in line:
            t_finish = now
        if now - t0 > timeout:
            p.kill()
            break
    rc = p.wait()
    dt = time.perf_counter() - t0
    render = (t_finish - t_start) if (t_start is not None and t_finish is not None) else None
    return dt, rc, "".join(tail)[-400:], render


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "ref_gpu_compare.json"))
    ap.add_argument("--timeout", type=float, default=240.0)
    a = ap.parse_args()
    import __graft_entry__ as ge
    rt = ge.load_package()

    # (scene file, width, height, reference spp pair, our spp)
    cases = [
        ("rtiow_486.json", 1920, 1080, (4, 16), 64),
        ("blue2.json", 1920, 1080, (64, 256), 256),
        ("sample_scene.json", 1600, 900, (128, 512), 512),
    ]
    results = []
    for name, W, H, (s1, s2), ours_spp in cases:
        scene = json.load(open(os.path.join(GOLDEN, name)))
        scene["width"], scene["height"] = W, H
        ref_scene = reference_schema(scene)
        rec = {"scene": name, "width": W, "height": H, "objects": len(scene["object"]["data"])}
        td = tempfile.mkdtemp(prefix="refgpu_")
        try:
            os.makedirs(os.path.join(td, "build", "output"), exist_ok=True)
            walls = {}
            for spp in (s1, s2):
                ref_scene["samples_per_pixel"] = spp
                ref_scene["output_file"] = "./build/output/ref.png"
                fn = os.path.join(td, f"ref_{spp}.json")
                json.dump(ref_scene, open(fn, "w"))
                if os.path.exists(REF_BIN):
                    dt, rc, err, render = run_wall([REF_BIN, "-f", fn], td, a.timeout)
                    walls[spp] = render
                    rec.setdefault("reference_runs", []).append({"spp": spp, "wall_s": dt, "render_s": render, "exit": rc,
                                                                 "stderr_tail": err if rc else ""})
            if walls.get(s2) and all(r["exit"] == 0 for r in rec["reference_runs"]):
                rec["reference_render_s_per_spp"] = walls[s2] / s2    # the larger run
                rec["reference_samples_per_s"] = W * H * s2 / walls[s2]
            # ours, resident scene, same file (reference schema: same objects, constant background)
            ref_scene["samples_per_pixel"] = ours_spp
            fn = os.path.join(td, "ours.json")
            json.dump(ref_scene, open(fn, "w"))
            r = rt.Renderer(0)
            r.upload_json(fn)
            r.render(seed=1); r.synchronize()                         # warm-up (allocations, first launches)
            best = 1e30
            for _ in range(3):
                t0 = time.perf_counter()
                r.render(seed=1); r.synchronize()
                best = min(best, time.perf_counter() - t0)
            _, _, stats = r.download()
            r.close()
            rec["ours_spp"] = ours_spp
            rec["ours_render_s_per_spp"] = best / ours_spp
            rec["ours_samples_per_s"] = W * H * ours_spp / best
            rec["ours_rays"] = stats.get("rays")
            # whole-process, like for like with the reference's larger run
            ref_scene["samples_per_pixel"] = s2
            ref_scene["output_file"] = "./build/output/ours.png"
            fn2 = os.path.join(td, f"ours_{s2}.json")
            json.dump(ref_scene, open(fn2, "w"))
            if os.path.exists(OUR_CLI):
                dt, rc, err, _ = run_wall([OUR_CLI, "-f", fn2], td, a.timeout)
                rec["ours_cli_wall_s"] = {"spp": s2, "wall_s": dt, "exit": rc}
            if rec.get("reference_samples_per_s"):
                rec["speedup_render"] = rec["ours_samples_per_s"] / rec["reference_samples_per_s"]
        except subprocess.TimeoutExpired as e:
            rec["error"] = f"timeout after {a.timeout}s: {e.cmd[0]}"
        finally:
            shutil.rmtree(td, ignore_errors=True)
        print(json.dumps(rec), flush=True)
        results.append(rec)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump({"what": "reference gpu-version/main.cu (sm_100 build, unmodified) vs librt_b200 on the same B200", "cases": results},
              open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
