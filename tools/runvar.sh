# A/B of traversal-TU builds (make -C qaray_b200/csrc variant NAME=x DEFS=...): VARIANTS="a b c" bash tools/runvar.sh
# copies each variants/lib_<v>.so over the product library, runs the bench without the CPU leg and prints one line per variant
cp qaray_b200/libqaray_b200.so /tmp/lib_keep.so
for v in $VARIANTS; do cp variants/lib_$v.so qaray_b200/libqaray_b200.so; python bench.py --no-cpu-baseline --strong-spp 0 --steps ${STEPS:-3} --warmup 3 ${BENCH_ARGS} > /tmp/o.json 2>/tmp/o.err; python - "$v" <<'PY'
import json,sys
try:
    d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1]); r=d['roofline']['class_ms_per_step']
    print(sys.argv[1], round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'trace', round(r['trace'],2), 'shadow', round(r['shadow'],2), 'shade', round(r['shade'],2))
except Exception as e:
    print(sys.argv[1], 'FAILED', open('/tmp/o.err').read()[-300:])
PY
done
cp /tmp/lib_keep.so qaray_b200/libqaray_b200.so
