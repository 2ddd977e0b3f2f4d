cd /root/repo
for wl in large_pyramid many_pyramids tumbler joint_grid; do
  steps=120; [ $wl = tumbler ] && steps=30
  timeout 900 python bench.py --workload $wl --steps 200 --warmup 60 --cpu-steps $steps 2>&1 | tail -1 > gpurun_out/table_$wl.json
done
timeout 600 python bench.py --workload batched_pyramids --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 > gpurun_out/table_batched.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 2>&1 | tail -1 > gpurun_out/table_reference.json
python - <<'PY'
import json
for wl in ("large_pyramid","many_pyramids","tumbler","joint_grid","batched"):
    try:
        d=json.load(open(f'gpurun_out/table_{wl}.json'))
    except Exception as e:
        print(wl, 'FAILED', open(f'gpurun_out/table_{wl}.json').read()[-300:]); continue
    cb=d.get('cpu_baseline',{})
    print(wl, 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,2), 'l2res', round(d['value_l2_resident']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'cpu', round(cb.get('value',0)/1e6,3), cb.get('sample'), 'roof', round(d['roofline']['frac'],4), round(d['roofline']['achieved'],1), {k:round(v,4) for k,v in d['phase_ms'].items()}, d['config'].get('bodies_per_gpu'), d['config'].get('touching_contacts_per_gpu'), d['clocks'])
print(open('gpurun_out/table_reference.json').read()[:400])
PY
