set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r2n_tests.txt 2>&1; tail -5 gpurun_out/r2n_tests.txt
python bench.py > gpurun_out/r2n_bench_c2.json 2> gpurun_out/r2n_bench_c2.err; tail -c 3000 gpurun_out/r2n_bench_c2.json; tail -3 gpurun_out/r2n_bench_c2.err
for c in C3 C5; do python bench.py --config $c --no-e2e --no-cpu > gpurun_out/r2n_bench_$c.json 2> gpurun_out/r2n_bench_$c.err; tail -c 1500 gpurun_out/r2n_bench_$c.json; done
