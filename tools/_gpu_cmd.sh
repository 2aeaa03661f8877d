timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^\[b200\]" | tail -5
timeout 100 python bench.py --workload C2-noshade --steps 15 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'frac', round(d['roofline']['frac'],3))"
