run() { python bench.py --no-cpu-baseline --steps 5 --warmup 3 "$@" > /tmp/o.json 2>/tmp/o.err
python - "NOPRIO=$QB_NO_PRIORITY $*" <<'PY' >> gpurun_out/ab.txt
import json,sys
try:
    d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1]); r=d['roofline']['class_ms_per_step']
    print(sys.argv[1], round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'trace', round(r['trace'],2), 'shadow', round(r['shadow'],2), 'shade', round(r['shade'],2), 'launches', d['gpu_launches'])
except Exception as e:
    print(sys.argv[1], 'FAILED', open('/tmp/o.err').read()[-300:])
PY
}
run --scene trc_scene_tower
run --scene trc_scene_xmas
export QB_NO_PRIORITY=1
run --scene trc_scene_tower
run --scene trc_scene_xmas
unset QB_NO_PRIORITY
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> gpurun_out/ab.txt
cat gpurun_out/ab.txt
