#!/bin/bash
# Closing measurement set of round 2 on one GPU box (each item writes under gpurun_out/).
mkdir -p gpurun_out
# 1. converged renders against the reference's own images (PSNR lines)
timeout 600 python -m pytest tests -m gpu -x -q -s -k converged > gpurun_out/psnr.log 2>&1; grep "psnr=" gpurun_out/psnr.log
# 2. checked build (device asserts on every queue / staging / stack index; compute-sanitizer is closed on this pool)
( echo "# RT_CHECKED=1 build (tools/build_variant.py checked -DRT_CHECKED=1): device-side asserts; $(date -u +%F)"; 
  RT_B200_LIB=build_variants/checked.so timeout 600 python tools/sanitize_run.py 2>&1 | tail -12;
  RT_B200_LIB=build_variants/checked.so timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_parity.py::test_cli_matches_the_library_and_the_reference_contract 2>&1 | tail -3 ) > gpurun_out/checked_build.log 2>&1; tail -4 gpurun_out/checked_build.log
# 3. the reference's own GPU program on this box vs the library (small scenes), and the queue-free schedule (tail from bounce 1)
timeout 400 python tools/ref_gpu_compare.py > gpurun_out/ref_gpu_compare.log 2>&1
python -c "
import json
for c in json.load(open('gpurun_out/ref_gpu_compare.json'))['cases']:
    print(c['scene'], 'reference GPU Msamples/s', round((c.get('reference_samples_per_s') or 0)/1e6,1), 'ours', round(c['ours_samples_per_s']/1e6,1), 'x', round(c.get('speedup_render',0),1))
"
for tb in 8 1 2; do for wl in blue2_1080p_256spp sample_scene_1600x900_512spp; do echo "tail_bounce=$tb $wl $(RT_TAIL_BOUNCE=$tb python tools/stage_bench.py $wl 128 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})")"; done; done | tee gpurun_out/queue_free.log
# 4. reference arm of the bench
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 400 gpurun_out/bench_reference.json
# 5. mesh: LBVH build time + ncu of the extend launches
python - <<'PY' | tee gpurun_out/mesh_build.log
import sys; sys.path.insert(0, '.')
import __graft_entry__ as ge
rt = ge.load_package()
from rt_b200 import scenes
r = rt.Renderer(0)
ts = scenes.mesh_scene(708, 708, 1920, 1080, 16)
for k in range(3):
    r.upload(ts); r.render(seed=1); _, _, st = r.download(want_sum=False, want_rgb8=False)
    b = r.bvh()
    print("1 002 532 objects: ms_build", round(st["ms_build"], 3), "n_tree", b["n_tree"], "giants", b["n_giants"], "depth", b["depth"])
r.close()
PY
ncu --set full --clock-control none --import-source on -k regex:'k_extend' -c 3 -o gpurun_out/prof_mesh_extend -f python tools/profile_run.py mesh1m_1080p_64spp 1 8 > gpurun_out/ncu_mesh.log 2>&1
ls gpurun_out | tr '\n' ' '
