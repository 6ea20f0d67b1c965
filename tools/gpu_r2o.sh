#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_generator_gpu.py -m gpu -q -x 2>&1 | tail -n 3
for i in 1 2 3; do timeout 600 python bench.py --no-cpu-baseline --no-train 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), round(d['roofline']['frac'],3), 'norm', round(d['roofline_norm']['frac'],3), round(d['roofline_norm']['achieved']), d['clocks']['sm_mhz'], round(d['parity_check']['psnr_db'],2))"; done
