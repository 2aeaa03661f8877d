mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "sort_last" 2>&1 | grep "sort-last"
