# round-2 final pass (one gpurun call, one GPU): smoke, the GPU suite, the bench lines, the reference arm, ncu of the segment kernels
set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3b_smoke.txt 2>&1; tail -2 gpurun_out/r3b_smoke.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r3b_gpu.txt 2>&1; tail -2 gpurun_out/r3b_gpu.txt
python bench.py --impl reference > gpurun_out/r3b_bench_ref.json 2> gpurun_out/r3b_bench_ref.err; tail -c 600 gpurun_out/r3b_bench_ref.json
python bench.py > gpurun_out/r3b_bench.json 2> gpurun_out/r3b_bench.err; tail -c 1200 gpurun_out/r3b_bench.json; tail -2 gpurun_out/r3b_bench.err
for c in C3 C4 C5; do python bench.py --config $c > gpurun_out/r3b_bench_$c.json 2> gpurun_out/r3b_bench_$c.err; tail -c 300 gpurun_out/r3b_bench_$c.json; done
python scripts/bench_deflate.py > gpurun_out/r3b_deflate.json 2> gpurun_out/r3b_deflate.err; tail -c 400 gpurun_out/r3b_deflate.json
C4="python bench.py --config C4 --no-e2e --no-cpu --steps 1 --warmup 1"
ncu --set full --clock-control none --import-source on -k regex:"k_seg_find|k_lane_seg|k_seg_lz|k_seg_sync" -s 4 -c 5 -f -o gpurun_out/r3b_prof_seg $C4 > gpurun_out/r3b_ncu_seg.log 2>&1
python scripts/ncu_summary.py gpurun_out/r3b_prof_seg.ncu-rep > gpurun_out/r3b_ncu_seg_summary.txt; head -16 gpurun_out/r3b_ncu_seg_summary.txt | cut -c1-160
rm -f gpurun_out/r3b_prof_seg.ncu-rep
