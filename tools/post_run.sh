set -x
timeout 900 python -m pytest tests/test_decode.py tests/test_gpu_postprocess.py tests/test_gpu_graspnet.py -x -q -m gpu 2>&1 | tail -15
timeout 600 python bench.py --steps 20 --warmup 3 --no-training > gpurun_out/bench_post.json 2> gpurun_out/bench_post.err; tail -3 gpurun_out/bench_post.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_post.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','latency','gpu_launches')}, d['e2e']['value'])
PY
