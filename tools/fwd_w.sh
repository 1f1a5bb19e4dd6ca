TAG=${1:-w}; mkdir -p gpurun_out
{
for bs in 2 3 4; do for as in 4 5; do echo "a $as b $bs: $(CPOLY_TC_STAGES=$as CPOLY_TC_BSTAGES=$bs timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 0.05 2>&1 | tail -1)"; done; done
} > gpurun_out/${TAG}_w.txt 2>&1; cat gpurun_out/${TAG}_w.txt
CPOLY_TC_BSTAGES=4 CPOLY_TC_TRACE=3 timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 0.05 2> gpurun_out/${TAG}_trace.txt | tail -1
