set -x
timeout 900 python -m pytest tests/test_gpu_seg.py tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r2x_tests.txt 2>&1; tail -3 gpurun_out/r2x_tests.txt
for c in C2 C3 C4; do timeout 900 python bench.py --config $c --no-e2e --no-cpu --no-strong > gpurun_out/r2x_bench_$c.json 2> gpurun_out/r2x_bench_$c.err; tail -c 200 gpurun_out/r2x_bench_$c.json; done
