set -x
ZSB_TRACE=1 timeout 900 python -m pytest tests/test_gpu_seg.py -x -q -m gpu -s > gpurun_out/r3k_seg.txt 2>&1; tail -3 gpurun_out/r3k_seg.txt; grep "wave(s)" gpurun_out/r3k_seg.txt | sort | uniq -c | sort -rn | head -8
