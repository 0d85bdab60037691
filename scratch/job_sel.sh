set -x
timeout 900 python -m pytest tests/test_gpu_stages.py -x -q -m gpu -k "select" 2>&1 | tail -5
timeout 600 python scratch/sel_scatter.py 2>&1 | tail -5
