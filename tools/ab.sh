source tools/variants.sh true
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
run l96
run l64 -DEPLAN_LSU_MIN=64
run l128 -DEPLAN_LSU_MIN=128
run l192 -DEPLAN_LSU_MIN=192
run ltma -DEPLAN_LSU_MIN=100000
