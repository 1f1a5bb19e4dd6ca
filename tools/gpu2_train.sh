mkdir -p gpurun_out
nproc; cat /proc/cpuinfo | grep "model name" | head -1; uptime
for i in 1 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$i bench.py --gpus 2 --skip fp32,dcn_sweep,cpu,sustained,idd_polar --no-cpu-baseline 2>/dev/null | grep "^{" | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); t=d['train']; print('N=2 train', t['value'], t['ms_per_step'], t['allreduce']['exposed_ms_per_step'], 'infer', d['value'])"
done
timeout 600 python bench.py --skip fp32,dcn_sweep,cpu,sustained,idd_polar --no-cpu-baseline 2>/dev/null | grep "^{" | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); t=d['train']; print('N=1 train', t['value'], t['ms_per_step'])"
