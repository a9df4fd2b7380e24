for i in 1 2 3; do python -m pytest tests/test_gpu_fxaa_shard_bvh.py tests/test_gpu_parity.py -x -q 2>&1 | grep -E "^E |passed|failed|Error|^tests.*:[0-9]+:" | head -30; done
