mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dcn.py -q -x -p no:cacheprovider -k "forward or random_shapes or fused or module" 2>&1 | tail -3
for cfg in "256 256 64 128 1" "256 128 64 128 1" "256 64 64 128 1" "512 256 32 64 1" "128 64 128 256 1"; do echo "cfg $cfg: $(timeout 120 python tools/fwd_one.py $cfg bf16 1 0.05 2>&1 | tail -1)"; done
bash tools/frame_ab.sh X=1 | tail -1
