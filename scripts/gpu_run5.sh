set -x
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/r5_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r5_smoke.log
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/r5_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r5_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --members 65536 --distinct 512 --lanes 0 --no-e2e --no-cpu > gpurun_out/r5_bench_l0.json 2> gpurun_out/r5_bench_l0.err; echo "bench rc=$?"; python -c "import json;j=json.load(open('gpurun_out/r5_bench_l0.json'));print(j['value'],j['ms_per_step'],j['roofline']['frac'])"; tail -2 gpurun_out/r5_bench_l0.err
CMD="python bench.py --steps 1 --warmup 3 --members 65536 --distinct 512 --no-e2e --no-cpu --lanes 0"
ncu --set full --clock-control none --import-source on -k regex:"k_lane|k_lz" -s 6 -c 2 -o gpurun_out/r5_prof_lane $CMD > gpurun_out/r5_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r5_ncu.log
