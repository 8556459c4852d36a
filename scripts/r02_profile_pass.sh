# round-2 profile pass (one gpurun call): overlap experiment, C2 launch list, ncu --set full of k_check / k_big on C4
set -x
python scripts/overlap_experiment.py > gpurun_out/r2k_overlap.json 2> gpurun_out/r2k_overlap.err; tail -c 900 gpurun_out/r2k_overlap.json; tail -3 gpurun_out/r2k_overlap.err
C2="python bench.py --no-e2e --no-strong --no-cpu --steps 4 --warmup 3"
$C2 > gpurun_out/r2k_plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 6 -c 8 --csv --log-file gpurun_out/r2k_launches.csv $C2 > gpurun_out/r2k_ncu_c2.log 2>&1
C4="python bench.py --config C4 --no-e2e --no-cpu --steps 1 --warmup 3"
$C4 > gpurun_out/r2k_plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_check|k_big" -s 6 -c 2 -f -o gpurun_out/r2k_prof_c4 $C4 > gpurun_out/r2k_ncu_c4.log 2>&1
python scripts/ncu_summary.py gpurun_out/r2k_prof_c4.ncu-rep > gpurun_out/r2k_ncu_c4_summary.txt
tail -2 gpurun_out/r2k_ncu_c4.log; cat gpurun_out/r2k_ncu_c4_summary.txt; cat gpurun_out/r2k_launches.csv | tail -12; ls -la gpurun_out/r2k*
