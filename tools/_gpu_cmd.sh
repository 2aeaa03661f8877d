mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 tools/multigpu_check.py 2>&1 | grep "^{" | tail -2
for EX in p2p nccl; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $NG --workload C5-small --exchange $EX --steps 20 --warmup 5 > gpurun_out/r2c_sl_$EX.json 2> gpurun_out/r2c_sl_$EX.err; echo "sort-last $EX rc=$?"
tail -1 gpurun_out/r2c_sl_$EX.json | python -c "
import sys, json
d=json.loads(sys.stdin.read()); print('  ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'value', round(d['value']), 'launches', d['gpu_launches'])" || tail -5 gpurun_out/r2c_sl_$EX.err
done
