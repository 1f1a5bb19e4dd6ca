# forward-kernel iteration check (one gpurun call): forward parity tests, then the 64->64 layer and a few others
TAG=${1:-it}; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dcn.py -q -x -p no:cacheprovider -k "forward or fwd or module or fused" 2>&1 | tail -8 > gpurun_out/${TAG}_tests.txt; cat gpurun_out/${TAG}_tests.txt
{
for std in 2.0 0.05; do for as in 4 5; do echo "std $std a_stages $as: $(CPOLY_TC_STAGES=$as timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 $std 2>&1 | tail -1)"; done; done
for cfg in "128 128 128 256 1 bf16 1" "256 256 64 128 1 bf16 1" "512 256 32 64 1 bf16 1" "64 64 256 512 1 fp32 1" "256 256 64 128 16 bf16 1"; do
  echo "cfg $cfg: $(timeout 120 python tools/fwd_one.py $cfg 2>&1 | tail -1)"
done
} > gpurun_out/${TAG}_fwd.txt 2>&1; cat gpurun_out/${TAG}_fwd.txt
CPOLY_TC_TRACE=3 timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 2> gpurun_out/${TAG}_trace.txt | tail -1
