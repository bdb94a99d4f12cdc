set -x
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_j.log 2>&1; tail -2 gpurun_out/smoke_j.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_j.log 2>&1; tail -2 gpurun_out/pytest_j.log
python bench.py > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_j.json 2> gpurun_out/bench_ref_j.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_j.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"raster_kernel|setup_kernel|gt_norm_kernel|finalize_live_kernel" -s 16 -c 4 -o gpurun_out/stage_j -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:"setup_kernel|raster_kernel|gt_norm_kernel|finalize_live_kernel" -s 24 -c 48 --csv --log-file gpurun_out/traffic_j.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
cat gpurun_out/bench_j.json | cut -c1-400
