# usage: update_profiles.sh TAG — copy the judged summaries of GPU run TAG from gpurun_out/ (scratch) into profiles/ (tracked)
TAG=$1; R=r01
cp gpurun_out/${TAG}_bench.json profiles/${R}_bench.json
cp gpurun_out/${TAG}_bench_ref.json profiles/${R}_bench_ref.json
cp gpurun_out/${TAG}_gpu.txt profiles/${R}_gpu.txt
cp gpurun_out/${TAG}_launches.csv profiles/${R}_launches.csv
python scripts/ncu_summary.py gpurun_out/${TAG}_prof.ncu-rep > profiles/${R}_ncu_summary.txt
echo "source run: ${TAG}" >> profiles/${R}_ncu_summary.txt
