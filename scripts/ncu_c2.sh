# ncu --set full capture of the C2 step's kernels (k_lane, k_lz) for the product and a variant build.
# usage (under gpurun): bash scripts/ncu_c2.sh TAG [variant suffix ...]
TAG=$1; shift
CMD="python bench.py --no-e2e --no-strong --no-cpu --steps 1 --warmup 3"
for v in "" "$@"; do
  export ZSB_LIB=$PWD/zlib-streams_b200/libzsb$v.so
  $CMD > gpurun_out/${TAG}_plain$v.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"k_lane|k_lz" -s 6 -c 2 -f -o gpurun_out/${TAG}_prof$v $CMD > gpurun_out/${TAG}_ncu$v.log 2>&1
  echo "variant '$v' rc=$?"; tail -2 gpurun_out/${TAG}_ncu$v.log
done
ls -la gpurun_out/*.ncu-rep
