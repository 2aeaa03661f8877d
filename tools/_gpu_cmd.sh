timeout 300 python -m pytest tests -m gpu -x -q -s -k "edge or transparent" 2>&1 | grep -v "^\[b200\]" | tail -14
