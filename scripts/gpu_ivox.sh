# iVox-mode GPU tests + whole GPU suite + quick bench
timeout 600 python -m pytest tests/test_gpu_ivox.py -x -q 2>&1 | tail -25
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_ivox.py 2>&1 | tail -5
bash scripts/gpu_quick.sh "$1"
