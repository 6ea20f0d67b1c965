# same-box A/B of mat_norm_kernel chunking through bench.py (usage under gpurun: bash tools/norm_chunk_ab.sh)
for i in 1 2; do
  for v in 16 32 64; do
    S2P_NORM_PIX_PER_THREAD=$v timeout 200 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('ppt=$v', round(d['value']), 'norm', round(d['roofline_norm']['frac'],3), round(d['roofline_norm']['avg_launch_ms']*1e3,1),'us', 'clk', d['clocks']['sm_mhz'])
"
  done
done
