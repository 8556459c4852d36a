# one gpurun call: the no-allocation test alone in a fresh process, then the whole GPU suite (it runs last there too)
set -x
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2l_noalloc_alone.txt 2>&1; tail -5 gpurun_out/r2l_noalloc_alone.txt
python -m pytest tests -x -q -m gpu > gpurun_out/r2l_gpu.txt 2>&1; tail -8 gpurun_out/r2l_gpu.txt
