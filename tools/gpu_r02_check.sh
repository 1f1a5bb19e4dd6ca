# one gpurun call: GPU tests (all, no -x), smoke, the bench line.  Usage: gpurun -- bash tools/gpu_r02_check.sh <tag> [bench flags]
TAG=${1:-r02}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 1700 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -60 > gpurun_out/${TAG}_tests.txt; tail -15 gpurun_out/${TAG}_tests.txt
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -3 > gpurun_out/${TAG}_smoke.txt; cat gpurun_out/${TAG}_smoke.txt
timeout 900 python bench.py "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -5 gpurun_out/${TAG}_bench.err; cat gpurun_out/${TAG}_bench.json
