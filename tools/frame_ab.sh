# A/B of the inference frame with an environment knob: bash tools/frame_ab.sh "<ENV=val>" 
mkdir -p gpurun_out
for env in "$@"; do
  echo "== $env: $(env $env timeout 300 python bench.py --skip fp32,train,idd_polar,dcn_sweep,cpu,sustained --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])")"
done
