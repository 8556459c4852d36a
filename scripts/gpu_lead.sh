# usage: gpu_lead.sh TAG — e2e leg of bench.py with 0, 1, 2, 4, 6 leading chunks through k_big (diagnostic sweep)
TAG=$1
mkdir -p gpurun_out
for L in 0 2 4 6 1 0; do
  timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu --lead-big $L > gpurun_out/${TAG}_lead$L.json 2> gpurun_out/${TAG}_lead$L.err
  python -c "
import json;j=json.load(open('gpurun_out/${TAG}_lead$L.json'));print('lead',$L,'e2e',round(j['e2e']['value'],2),'GB/s',round(j['e2e']['ms_per_step'],2),'ms; value',round(j['value'],1))"
done
ZSB_TRACE=1 timeout 400 python bench.py --steps 3 --warmup 3 --no-cpu --lead-big 4 > /dev/null 2> gpurun_out/${TAG}_trace4.err; grep "zsb trace" gpurun_out/${TAG}_trace4.err | tail -34 | head -12
