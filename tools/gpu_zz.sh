# the allocation-free entry point alone in a fresh process, then the whole round-end sequence
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/zz_alone.txt; cat gpurun_out/zz_alone.txt
bash tools/gpu_r02_check.sh "$@"
