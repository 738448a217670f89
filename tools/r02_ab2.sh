for v in "$@"; do
 for wl in rtiow_1080p_1024spp sample_scene_1600x900_512spp blue2_1080p_256spp; do echo "== [$v] $wl"; env $v python tools/stage_bench.py $wl 128 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})"; done; done
