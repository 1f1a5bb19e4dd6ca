TAG=${1:-abl}; mkdir -p gpurun_out
for d in 0 7 4095; do CPOLY_TC_TRACE=3 CPOLY_TC_DBG=$d timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 2> gpurun_out/${TAG}_trace_$d.txt | tail -1; grep -c TRACE gpurun_out/${TAG}_trace_$d.txt; done
