for v in "" _vB _vC _vD; do
  ZSB_LIB=$PWD/zlib-streams_b200/libzsb$v.so python bench.py --no-e2e --no-strong --no-cpu --steps 10 > gpurun_out/r2d_bench$v.json 2> gpurun_out/r2d_bench$v.err
  python - <<PY
import json
j=json.load(open("gpurun_out/r2d_bench$v.json"))
print("variant '$v'", round(j["value"],1), "GB/s", round(j["ms_per_step"],2), "ms", j["roofline"]["kernel_ms_split"])
PY
done
python -m pytest tests/test_gpu_aux_api.py -q -m gpu 2>&1 | grep -E "reference != product|passed|failed" | cut -c1-1200
