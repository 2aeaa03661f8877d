mkdir -p gpurun_out
nvidia-smi -L | wc -l
NG=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 tools/multigpu_check.py 2>&1 | grep "^{" | tail -2
for N in 1 2 4 8; do
if [ $N -le $NG ]; then
if [ $N = 1 ]; then timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r1z_sf_$N.json 2> gpurun_out/r1z_sf_$N.err;
else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r1z_sf_$N.json 2> gpurun_out/r1z_sf_$N.err; fi
echo "sort-first C2 N=$N rc=$?"; tail -1 gpurun_out/r1z_sf_$N.json | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('  ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'value', round(d['value']), 'launches', d['gpu_launches'])"
fi
done
for N in 2 8; do
if [ $N -le $NG ]; then
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --workload C3 --steps 10 --warmup 3 > gpurun_out/r1z_sf_C3_$N.json 2> gpurun_out/r1z_sf_C3_$N.err
echo "sort-first C3 N=$N rc=$?"; tail -1 gpurun_out/r1z_sf_C3_$N.json | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('  ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'value', round(d['value']))"
fi
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $NG --workload C5 --steps 5 --warmup 2 > gpurun_out/r1z_sl_C5_$NG.json 2> gpurun_out/r1z_sl_C5_$NG.err
echo "sort-last C5 N=$NG rc=$?"; tail -1 gpurun_out/r1z_sl_C5_$NG.json | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('  ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'value', round(d['value']), d['config']['parallelism'])"
tail -3 gpurun_out/r1z_sl_C5_$NG.err
