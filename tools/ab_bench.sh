show() { python -c "
import sys,json
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']
print(sys.argv[1], d['value'], d['ms_per_step'], {n:round(v,2) for n,v in k.items() if 'lstm_tc' in n}, d['clocks'])" $1; }
cp masters-thesis_b200/libb200lstm.so /tmp/new.so
for i in 1 2; do
timeout 40 python bench.py --kernels-only --steps 5 --warmup 3 > /tmp/new$i.json 2>/dev/null; show /tmp/new$i.json
cp masters-thesis_b200/build/alt_head.so masters-thesis_b200/libb200lstm.so
timeout 40 python bench.py --kernels-only --steps 5 --warmup 3 > /tmp/old$i.json 2>/dev/null; show /tmp/old$i.json
cp /tmp/new.so masters-thesis_b200/libb200lstm.so
done
