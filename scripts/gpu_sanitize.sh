# usage: gpu_sanitize.sh TAG — compute-sanitizer over the small-batch GPU parity tests (memcheck / racecheck
# on the decode kernels).  memcheck: every decoder path over the golden fixtures, the probe vectors, the damaged and
# truncated streams of Deflate64, the ZIP archive and the z_stream surface; racecheck + synccheck: the shared-memory
# users (k_lane rings and tables, k_big queue, k_lz).  Logs land in gpurun_out/${TAG}_*.log.
TAG=$1
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
MEM_TESTS="tests/test_gpu_parity.py::test_golden_fixtures tests/test_gpu_parity.py::test_probe_vectors tests/test_gpu_parity.py::test_many_small_and_unaligned tests/test_gpu_parity.py::test_call_level_edge_cases tests/test_gpu_parity.py::test_zstream_wasm_pattern tests/test_gpu_parity.py::test_zstream_many_streams_per_call tests/test_gpu_parity.py::test_host_batch_pipelined tests/test_deflate64.py::test_gpu_deflate64 tests/test_zip.py"
RACE_TESTS="tests/test_gpu_parity.py::test_golden_fixtures tests/test_gpu_parity.py::test_probe_vectors tests/test_deflate64.py::test_gpu_deflate64"
timeout 1500 $CS --tool memcheck --error-exitcode 9 --print-limit 20 --log-file gpurun_out/${TAG}_memcheck.log \
    python -m pytest $MEM_TESTS -m gpu -x -q > gpurun_out/${TAG}_memcheck_pytest.log 2>&1
echo "memcheck rc=$?"; tail -2 gpurun_out/${TAG}_memcheck_pytest.log | cut -c1-200; tail -3 gpurun_out/${TAG}_memcheck.log
timeout 1500 $CS --tool racecheck --racecheck-report all --error-exitcode 9 --print-limit 20 --log-file gpurun_out/${TAG}_racecheck.log \
    python -m pytest $RACE_TESTS -m gpu -x -q > gpurun_out/${TAG}_racecheck_pytest.log 2>&1
echo "racecheck rc=$?"; tail -2 gpurun_out/${TAG}_racecheck_pytest.log | cut -c1-200; tail -3 gpurun_out/${TAG}_racecheck.log
timeout 900 $CS --tool synccheck --error-exitcode 9 --print-limit 20 --log-file gpurun_out/${TAG}_synccheck.log \
    python -m pytest $RACE_TESTS -m gpu -x -q > gpurun_out/${TAG}_synccheck_pytest.log 2>&1
echo "synccheck rc=$?"; tail -2 gpurun_out/${TAG}_synccheck_pytest.log | cut -c1-200; tail -3 gpurun_out/${TAG}_synccheck.log
