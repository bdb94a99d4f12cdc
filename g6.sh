P=/root/repo/end-to-end-rgb-pose-estimation-baseline_b200
for v in librac_b200.so librac_b200_sp9.so librac_b200_sp12.so librac_b200_sp20.so; do
  RAC_LIB=$P/$v timeout 200 python -m pytest tests/test_gpu_raster.py -x -q -k "matches_oracle or golden" 2>&1 | tail -1
  for i in 1 2; do RAC_LIB=$P/$v timeout 120 python bench.py --no-cpu-baseline --steps 100 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['roofline']['stage_ms'])"; done
done
