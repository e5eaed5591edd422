cd /root/repo
for i in $(seq 1 8); do
  timeout 300 python -m pytest tests/test_train_gpu.py -x -q -k "captured_step_equals_eager" 2>&1 | grep -E "AssertionError|passed|failed" | head -3
done
