#!/bin/bash
# nets tests + bench + phase timeline of the fused chain kernel (each under its own timeout)
timeout -k 5 60 python -m pytest tests/test_nets_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout -k 5 60 python bench.py --no-cpu-baseline --no-fp32-line > gpurun_out/bench_try.log 2>gpurun_out/bench_try.err
python -c "
import json;d=json.loads(open('gpurun_out/bench_try.log').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['stages'],d['roofline']['frac'],d['kept_grasps_last_step'])" || tail -3 gpurun_out/bench_try.err
rm -f gpurun_out/clocks_try.txt
DEXGG_CHAIN_CLOCKS=gpurun_out/clocks_try.txt timeout -k 5 40 python tools/chain_clocks.py > gpurun_out/clocks.log 2>&1
