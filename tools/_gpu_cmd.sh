for v in 3 4; do
  B200VR_NVCC_EXTRA="-DB200VR_Q_MIN_CTAS=$v" python -c "import ovr_b200; ovr_b200.capi.build(force=True)"
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1e_bench_$v.log 2>&1; tail -1 gpurun_out/r1e_bench_$v.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('min_ctas $v', d['value'], d['ms_per_step'], d['roofline']['frac'], d['config']['samples_per_frame'])"
done
timeout 900 python -m pytest tests -m gpu -q -s -x > gpurun_out/r1_gputests.log 2>&1; grep -E "passed|failed|FAILED|max-abs|Error|assert" gpurun_out/r1_gputests.log | tail -12
