set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r1_smi.txt 2>&1
nproc >> gpurun_out/r1_smi.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1; echo "smoke rc=$?" 
tail -5 gpurun_out/r1_smoke.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r1_pytest.log
for L in 32 8 4 1; do
  timeout 300 python bench.py --steps 3 --warmup 3 --members 16384 --distinct 256 --lanes $L --no-e2e --no-cpu > gpurun_out/r1_bench_l$L.json 2> gpurun_out/r1_bench_l$L.err; echo "bench lanes=$L rc=$?"; cat gpurun_out/r1_bench_l$L.json | cut -c1-400; tail -3 gpurun_out/r1_bench_l$L.err
done
