# usage: bash scripts/emul_sanitize.sh — every host-emulation test of the device decoder sources under ASan + UBSan
# (tests/emul/libemul_asan.so; CPU only, ~15 min).  Writes profiles/r01_emul_asan_ubsan.txt.
cd "$(dirname "$0")/.."
ASAN=$(gcc -print-file-name=libasan.so)
OUT=profiles/r01_emul_asan_ubsan.txt
{
  echo "# tests/emul (device .cuh sources, host build, -fsanitize=address,undefined -fno-sanitize-recover=undefined)"
  echo "# $(g++ --version | head -1); $(date -u +%Y-%m-%dT%H:%MZ)"
} > $OUT
for t in tests/test_emul_parity.py tests/test_emul_stream.py "tests/test_deflate64.py -k emul"; do
  log=$(mktemp)
  ZSB_EMUL_SANITIZE=1 LD_PRELOAD=$ASAN ASAN_OPTIONS=detect_leaks=0 timeout 3000 python -m pytest $t -q -s -p no:cacheprovider > $log 2>&1
  rc=$?
  echo "== $t: rc=$rc, $(grep -c 'runtime error' $log) UBSan reports, $(grep -c 'ERROR: AddressSanitizer' $log) ASan reports" >> $OUT
  tail -n 1 $log >> $OUT
  grep -m 5 -A 6 'runtime error\|ERROR: AddressSanitizer' $log | cut -c1-200 >> $OUT
  rm -f $log
done
cat $OUT
