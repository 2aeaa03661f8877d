mkdir -p gpurun_out
run() {
  B200VR_NVCC_EXTRA="$1" python -c "
import sys, importlib
sys.path.insert(0,'.')
m=importlib.import_module('open-volume-renderer_b200._capi'); m.build(force=True)" > /dev/null 2>&1 || echo "build failed $1"
  python bench.py --steps 15 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'frac', round(d['roofline']['frac'],3))"
}
run "-DB200VR_Q_THREADS=352 -DB200VR_Q_MIN_CTAS=2"
run "-DB200VR_Q_THREADS=224 -DB200VR_Q_MIN_CTAS=3"
run "-DB200VR_Q_THREADS=256 -DB200VR_Q_MIN_CTAS=3"
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
