set -x
timeout 900 python -m pytest tests/test_gpu_seg.py -x -q -m gpu > gpurun_out/r2q_seg.txt 2>&1; tail -5 gpurun_out/r2q_seg.txt
for c in C3 C4; do ZSB_TRACE=1 timeout 900 python bench.py --config $c --no-e2e --no-cpu > gpurun_out/r2q_bench_$c.json 2> gpurun_out/r2q_bench_$c.err; tail -c 400 gpurun_out/r2q_bench_$c.json; grep "zsb trace" gpurun_out/r2q_bench_$c.err | sort | uniq -c | head -4; done
for c in C3 C4; do
B="python bench.py --config $c --no-e2e --no-cpu --steps 1 --warmup 1"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_seg|k_lane|k_big|k_lz|k_check" -s 12 -c 22 --csv --log-file gpurun_out/r2q_launches_$c.csv $B > gpurun_out/r2q_ncu_$c.log 2>&1
done
