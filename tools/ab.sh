source tools/variants.sh true
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
run new
FEELGPU_EPLAN_NO_ALIGNED_S=1 run newgeneric
runref old
