# usage: gpu_quick.sh TAG [pytest]   — smoke, (parity tests), C2 bench on the lane path, ncu of k_lane/k_lz
TAG=$1
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
if [ "$2" = "pytest" ]; then timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log; fi
timeout 600 python bench.py --steps 3 --warmup 3 --members 65536 --distinct 512 --lanes 0 --no-e2e --no-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; python -c "import json;j=json.load(open('gpurun_out/${TAG}_bench.json'));print('GBps',j['value'],'ms',j['ms_per_step'],'frac',j['roofline']['frac'])"; tail -2 gpurun_out/${TAG}_bench.err
CMD="python bench.py --steps 1 --warmup 3 --members 65536 --distinct 512 --no-e2e --no-cpu --lanes 0"
ncu --set full --clock-control none --import-source on -k regex:"k_lane|k_lz" -s 6 -c 2 -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
