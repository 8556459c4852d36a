timeout 300 python scripts/bench_deflate.py > gpurun_out/r2h_deflate.json 2> gpurun_out/r2h_deflate.err; tail -c 1800 gpurun_out/r2h_deflate.json; tail -3 gpurun_out/r2h_deflate.err
timeout 500 python bench.py --config C4 --steps 3 --no-e2e > gpurun_out/r2h_C4.json 2>gpurun_out/r2h_C4.err
python -c "
import json;j=json.load(open('gpurun_out/r2h_C4.json'));print(j['value'],j['ms_per_step'],j['roofline']['kernel_ms_split'],j['roofline']['k_check_gbs'])"
