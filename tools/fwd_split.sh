mkdir -p gpurun_out
{
for cfg in "256 256 64 128 1" "256 128 64 128 1" "256 64 64 128 1" "128 128 128 256 1" "128 64 128 256 1" "512 256 32 64 1"; do
  for sp in 1 3 9; do echo "cfg $cfg split $sp: $(CPOLY_TC_SPLIT=$sp timeout 120 python tools/fwd_one.py $cfg bf16 1 0.05 2>&1 | tail -1)"; done
done
} > gpurun_out/split.txt 2>&1; cat gpurun_out/split.txt
