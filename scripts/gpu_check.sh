# GPU tests + bench summary (+ solve phase stamps)
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
bash scripts/gpu_quick.sh "$1"
