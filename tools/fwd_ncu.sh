# ncu (selected sections) of the forward kernel on the 64->64 @256x512 bf16 layer for one offset std (run AFTER the plain command exited 0)
TAG=${1:-ncu}; STD=${2:-2.0}; mkdir -p gpurun_out
timeout 120 python tools/fwd_one.py 64 64 256 512 1 bf16 1 $STD > gpurun_out/${TAG}_plain.txt 2>&1 || exit 1
cat gpurun_out/${TAG}_plain.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dcn_fwd_tc_kernel -s 3 -c 1 -o gpurun_out/${TAG}_fwd -f python tools/fwd_one.py 64 64 256 512 1 bf16 1 $STD > gpurun_out/${TAG}_ncu.log 2>&1
tail -2 gpurun_out/${TAG}_ncu.log
