NG=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 tools/multigpu_check.py 2>&1 | grep -v "^\[b200\]\|OMP_NUM\|^\*\*\*" | head -40
