P='import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["roofline"]["physics_kernel_ms_avg"], d["roofline"]["kernel_ms_avg"])'
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 | python -c "$P"
