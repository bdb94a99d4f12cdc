python -m pytest tests/test_gpu_raster.py -x -q -k "assemble or reference_api" 2>&1 | tail -3
python profiles/e2e_breakdown.py 2>&1 | tail -10
python bench.py --no-cpu-baseline --steps 30 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['stage_ms'], d['e2e'])"
