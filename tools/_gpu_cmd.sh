timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^\[b200\]" | tail -5
run() {
  env $1 timeout 60 python bench.py --steps 15 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"
}
run "B200VR_KERNEL=q"
run "B200VR_KERNEL=ring"
run "B200VR_KERNEL=ring B200VR_RING_HIGH=2"
run "B200VR_KERNEL=ring B200VR_RING_HIGH=8"
B200VR_KERNEL=ring timeout 100 python tools/partition_probe.py C2 2>&1 | grep -v "^\[b200\]"
