mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_batchnorm.py -q -x -p no:cacheprovider 2>&1 | tail -3
for env in "$@"; do
  echo "== $env: $(env $env timeout 400 python bench.py --skip fp32,idd_polar,dcn_sweep,cpu,sustained --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); t=d['train']; print(t['value'], t['ms_per_step'], t['loss_first'], t['loss_last'])")"
done
timeout 300 python tools/train_trace.py 2>&1 | grep -E "batchnorm|bn_|Self CUDA" | cut -c1-60,150-260 | head -12
