for i in $(seq 1 14); do timeout 600 python -m pytest tests/test_rasterizer_gpu.py -q --timeout 600 -x 2>&1 | grep -E "^E  |passed|failed|Error" | cut -c1-400 | head -12; done
