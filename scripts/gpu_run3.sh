set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r3_smi.txt 2>&1
nproc >> gpurun_out/r3_smi.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r3_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r3_smoke.log
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/r3_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r3_pytest.log
for L in 32 16 8 4 2 1; do
  timeout 300 python bench.py --steps 3 --warmup 3 --members 16384 --distinct 256 --lanes $L --no-e2e --no-cpu > gpurun_out/r3_bench_l$L.json 2> gpurun_out/r3_bench_l$L.err; echo "bench lanes=$L rc=$?"; python -c "import json;j=json.load(open('gpurun_out/r3_bench_l$L.json'));print(j['value'],j['ms_per_step'],j['roofline']['frac'])"; tail -2 gpurun_out/r3_bench_l$L.err
done
CMD="python bench.py --steps 1 --warmup 3 --members 8192 --distinct 64 --no-e2e --no-cpu --lanes 8"
$CMD > gpurun_out/r3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_inflate -s 3 -c 1 -o gpurun_out/r3_prof_l8 $CMD > gpurun_out/r3_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r3_ncu.log
