# usage: gpu_value.sh TAG — device-resident bench only (value, k_lane / k_lz times), no e2e, no CPU leg
TAG=$1
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/${TAG}_value.json 2> gpurun_out/${TAG}_value.err; echo "rc=$?"
python -c "
import json;j=json.load(open('gpurun_out/${TAG}_value.json'));print('${TAG} value',round(j['value'],1),'ms',round(j['ms_per_step'],2),'k_lane',round(j['roofline']['kernel_ms'],2),'k_lz',j['roofline']['other_kernels_ms'])"; tail -2 gpurun_out/${TAG}_value.err
