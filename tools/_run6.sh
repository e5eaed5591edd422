cd /root/repo
timeout 300 python tools/chain_time.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_nets_gpu.py tests/test_pipeline_gpu.py -x -q 2>&1 | tail -4
