# 2-GPU check: DP correctness test, then the bench under torchrun (N=2)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ddp.py -q -x -p no:cacheprovider 2>&1 | tail -6 > gpurun_out/r02m_ddp_2gpu_test.txt; cat gpurun_out/r02i_ddp_2gpu_test.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --skip fp32,dcn_sweep,cpu,sustained --no-cpu-baseline > gpurun_out/r02m_bench_2gpu.json 2> gpurun_out/r02m_bench_2gpu.err; tail -3 gpurun_out/r02m_bench_2gpu.err; grep "^{" gpurun_out/r02m_bench_2gpu.json | tail -1 | cut -c1-400
