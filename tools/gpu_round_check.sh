set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json
timeout 200 python tools/raster_bench.py > gpurun_out/raster_bench.jsonl 2>/dev/null; cat gpurun_out/raster_bench.jsonl
timeout 300 ncu --set full --clock-control none --import-source on -k regex:raster_ -c 3 -o gpurun_out/raster_b1 -f python tools/raster_ncu.py 1 > gpurun_out/ncu_raster.log 2>&1; tail -1 gpurun_out/ncu_raster.log
timeout 300 ncu --set full --clock-control none -k regex:raster_cover -c 1 -o gpurun_out/raster_b8 -f python tools/raster_ncu.py 8 > gpurun_out/ncu_raster8.log 2>&1; tail -1 gpurun_out/ncu_raster8.log
