# usage: gpu_e2e.sh TAG — GPU parity tests, then the default bench (device-resident + e2e + cpu baseline)
TAG=$1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${TAG}_pytest.log | cut -c1-400
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; python -c "
import json;j=json.load(open('gpurun_out/${TAG}_bench.json'));print('value',j['value'],'ms',j['ms_per_step'],'kernel_ms',j['roofline']['kernel_ms'],'lz',j['roofline']['other_kernels_ms'],'e2e',j['e2e']['value'],j['e2e']['ms_per_step'],'cpu',j['cpu_baseline']['value'])"; tail -3 gpurun_out/${TAG}_bench.err
