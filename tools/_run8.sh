cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/r02r_bench_n1.json 2> gpurun_out/r02r_bench_n1.err || tail -c 800 gpurun_out/r02r_bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02r_bench_n1.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["frac_of_sustained_peak"], d["parity_check"]["passed"], d["parity_check"]["max_dp"])
print(d["stages"]); print(d["clocks"]); print(d["train"]["ms_per_step"] if "train" in d else None, d["fp32_exact_mode"]["value"] if d.get("fp32_exact_mode") else None)
PY
