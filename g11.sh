python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -1
python bench.py > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_k.json 2> gpurun_out/bench_ref_k.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_k.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"raster_kernel|setup_kernel|gt_norm_kernel|finalize_live_kernel" -s 16 -c 4 -o gpurun_out/stage_k -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:"setup_kernel|raster_kernel|gt_norm_kernel|finalize_live_kernel" -s 24 -c 48 --csv --log-file gpurun_out/traffic_k.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
