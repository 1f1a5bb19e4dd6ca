TAG=${1:-w}; mkdir -p gpurun_out
{
for std in 0.05 2.0; do for pf in "-1 1" "1 1" "2 1" "3 1" "2 2" "6 1"; do set -- $pf; echo "std $std pf $1 dist $2: $(CPOLY_TC_PF=$1 CPOLY_TC_PFD=$2 timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 $std 2>&1 | tail -1)"; done; done
} > gpurun_out/${TAG}_w.txt 2>&1; cat gpurun_out/${TAG}_w.txt
