set -x
timeout 900 python -m pytest tests/test_gpu_seg.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r2w_tests.txt 2>&1; tail -3 gpurun_out/r2w_tests.txt
for v in "" _s1 _s4; do
  for c in C2 C3 C4; do
    ZSB_LIB=$PWD/zlib-streams_b200/libzsb$v.so timeout 900 python bench.py --config $c --no-e2e --no-cpu --no-strong > gpurun_out/r2w_bench_$c$v.json 2> gpurun_out/r2w_bench_$c$v.err
    python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/r2w_bench_$c$v.json").read().strip().splitlines()[-1])
    print("variant '$v' $c", round(j["value"],1), "GB/s", round(j["ms_per_step"],2), "ms", j["roofline"]["kernel_ms_split"])
except Exception as e:
    print("variant '$v' $c ERR", e)
PY
  done
done
