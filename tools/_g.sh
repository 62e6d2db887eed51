python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/layer_table.py 8 2>&1 | tail -27
