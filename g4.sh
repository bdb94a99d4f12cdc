python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python profiles/e2e_breakdown.py 2>&1 | tail -9
python bench.py --no-cpu-baseline --steps 50 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['stage_ms'], d['e2e'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_i.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
python profiles/summarize_launches.py gpurun_out/launches_i.csv | head -8
