for v in $VARIANTS; do cp variants/lib_$v.so qaray_b200/libqaray_b200.so; python bench.py --no-cpu-baseline --steps 3 --warmup 3 > /tmp/o.json 2>/tmp/o.err; python - "$v" <<'PY'
import json,sys
try:
    d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1]); r=d['roofline']['class_ms_per_step']
    print(sys.argv[1], round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'trace', round(r['trace'],2), 'shadow', round(r['shadow'],2), 'shade', round(r['shade'],2))
except Exception as e:
    print(sys.argv[1], 'FAILED', open('/tmp/o.err').read()[-300:])
PY
done
