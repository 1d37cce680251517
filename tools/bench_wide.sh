for eng in tc_fast tc_precise; do
timeout 250 python bench.py --model wide --engine $eng --batch 256 --steps 3 --warmup 3 --kernels-only 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
print('$eng', round(d['value'],1), 'samples/s', round(d['ms_per_step'],1), 'ms', 'achieved TF', round(d['roofline']['achieved'],1), {n:round(v,1) for n,v in k.items() if v>1}, d['roofline']['kernel_launches_per_step'].get('wide_gemm_rec_fwd'))"
done
