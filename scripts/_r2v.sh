set -x
timeout 900 python -m pytest tests/test_gpu_seg.py tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r2v_seg.txt 2>&1; tail -5 gpurun_out/r2v_seg.txt
for c in C3 C4 C5; do timeout 900 python bench.py --config $c --no-e2e --no-cpu > gpurun_out/r2v_bench_$c.json 2> gpurun_out/r2v_bench_$c.err; tail -c 300 gpurun_out/r2v_bench_$c.json; done
timeout 900 python bench.py --no-e2e --no-cpu --no-strong > gpurun_out/r2v_bench_C2.json 2> gpurun_out/r2v_bench_C2.err; tail -c 300 gpurun_out/r2v_bench_C2.json
