# per-kernel LSTM times of the bench workload: bash tools/kbench.sh [extra bench flags]
timeout 60 python bench.py --kernels-only --steps 5 --warmup 3 "$@" 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
print(round(d['value']), round(d['ms_per_step'],2), {n[8:]:round(v,2) for n,v in k.items() if 'lstm_tc' in n}, d['clocks']['sm_mhz'])"
