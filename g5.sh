timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do timeout 120 python bench.py --no-cpu-baseline --steps 100 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['stage_ms'], d['e2e']['ms_per_step'])"; done
