# timing experiment: backward-epilogue ablations (needs B200_ABLATE=1 build); results of these runs are numerically wrong by design
for m in 0 1 2 4 8 16 31; do
B200_TC_DBG=$m timeout 60 python bench.py --kernels-only --steps 3 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']
print('dbg=$m', {n[12:]:round(v,2) for n,v in k.items() if 'lstm_tc_bwd' in n})"
done
