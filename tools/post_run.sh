set -x
timeout 900 python -m pytest tests/test_gpu_postprocess.py tests/test_gpu_graspnet.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15
timeout 600 python bench.py --steps 20 --warmup 3 --no-training > gpurun_out/bench_post.json 2> gpurun_out/bench_post.err; tail -3 gpurun_out/bench_post.err; cat gpurun_out/bench_post.json
