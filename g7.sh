ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:"setup_kernel|raster_kernel|gt_norm_kernel|finalize_live_kernel" -s 24 -c 48 --csv --log-file gpurun_out/traffic_i.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/traffic_i.csv
