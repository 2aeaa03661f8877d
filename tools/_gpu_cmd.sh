set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1q_pre.json 2>&1; echo "pre rc=$?"
ncu --set full --clock-control none --import-source on -k regex:raymarch_q_kernel -s 3 -c 1 -f -o gpurun_out/r1q_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1q_ncu.log 2>&1; echo "ncu rc=$?"
