#!/bin/bash
# Round-3 evidence, one GPU (run on the box: gpurun -- 'bash profiles/collect_r03.sh').  Every ncu pass follows a plain run of
# the same command that exited 0; numbers printed under ncu are never bench values.
set -u
O=gpurun_out/r03
mkdir -p $O
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 200 > $O/clocks.csv &
SMI=$!
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py --steps 50 --warmup 5 > $O/bench.json 2> $O/bench.err || { echo "bench failed"; tail -5 $O/bench.err; }
kill $SMI
python profiles/e2e_breakdown.py --driver > $O/e2e_breakdown_driver.txt 2>/dev/null
python profiles/e2e_breakdown.py --u8 > $O/e2e_breakdown_u8.txt 2>/dev/null
python profiles/gt_norm_bench.py > $O/gt_norm_bench.txt 2>&1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
$CMD > $O/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:setup_kernel|raster_kernel|gt_norm_kernel|finalize_live_kernel' -s 40 -c 4 -o $O/stage $CMD > $O/ncu_full.log 2>&1
$CMD > $O/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k 'regex:setup_kernel|raster_kernel|gt_norm_kernel|finalize_live_kernel' -s 40 -c 48 --csv --log-file $O/traffic.csv $CMD > $O/ncu_traffic.log 2>&1
python profiles/gt_norm_bench.py child > $O/plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:gt_norm_tma_kernel|assemble_render_kernel' -s 20 -c 2 -o $O/assemble python profiles/gt_norm_bench.py child > $O/ncu_assemble.log 2>&1
ls -la $O
