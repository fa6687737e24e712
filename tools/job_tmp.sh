run() { python bench.py --no-cpu-baseline --steps 5 --warmup 3 "$@" > /tmp/o.json 2>/tmp/o.err
python - "SORT_RAYS=$QB_SORT_RAYS $*" <<'PY' >> gpurun_out/ab.txt
import json,sys
try:
    d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1]); r=d['roofline']['class_ms_per_step']
    print(sys.argv[1], round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'trace', round(r['trace'],2), 'shadow', round(r['shadow'],2), 'shade', round(r['shade'],2), 'launches', d['gpu_launches'])
except Exception as e:
    print(sys.argv[1], 'FAILED', open('/tmp/o.err').read()[-300:])
PY
}
for m in 0 1 2 3; do export QB_SORT_RAYS=$m; run --scene trc_scene_tower; done
export QB_SORT_RAYS=3
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> gpurun_out/ab.txt
run --scene trc_scene_xmas
cat gpurun_out/ab.txt
