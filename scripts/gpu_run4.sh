set -x
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/r4_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r4_smoke.log
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/r4_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r4_pytest.log
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_parity.py -x -q -k "probe_vectors and 0" > gpurun_out/r4_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -5 gpurun_out/r4_memcheck.log
for L in 0 16; do
  timeout 300 python bench.py --steps 3 --warmup 3 --members 16384 --distinct 256 --lanes $L --no-e2e --no-cpu > gpurun_out/r4_bench_l$L.json 2> gpurun_out/r4_bench_l$L.err; echo "bench lanes=$L rc=$?"; python -c "import json;j=json.load(open('gpurun_out/r4_bench_l$L.json'));print(j['value'],j['ms_per_step'],j['roofline']['frac'])"; tail -2 gpurun_out/r4_bench_l$L.err
done
CMD="python bench.py --steps 1 --warmup 3 --members 16384 --distinct 256 --no-e2e --no-cpu --lanes 0"
$CMD > gpurun_out/r4_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_lane|k_lz" -s 6 -c 2 -o gpurun_out/r4_prof_lane $CMD > gpurun_out/r4_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r4_ncu.log
