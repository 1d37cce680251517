# per-kernel times of the widened model (configs[2], bf16 engine): bash tools/wbench.sh [extra bench flags]
timeout 200 python bench.py --model wide --engine tc_fast --kernels-only --steps 2 --warmup 1 "$@" 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
print(round(d['value']), round(d['ms_per_step'],1), round(d['roofline']['frac'],3), {n:round(v,1) for n,v in k.items() if v>1.5}, d['clocks']['sm_mhz'])"
