timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^\[b200\]" | tail -12
