timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
for tb in 1 2 3; do echo "RT_TAIL_BOUNCE=$tb $(RT_TAIL_BOUNCE=$tb python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], d['image_check'])")"; done
