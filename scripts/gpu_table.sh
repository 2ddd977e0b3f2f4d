cd /root/repo
for wl in large_pyramid many_pyramids tumbler joint_grid; do
  steps=120; [ $wl = tumbler ] && steps=20
  timeout 600 python bench.py --workload $wl --steps 200 --warmup 60 --cpu-steps $steps 2>&1 | tail -1 > gpurun_out/table_$wl.json
  python -c "
import json,sys
d=json.load(open('gpurun_out/table_$wl.json'))
print(d['config']['workload'], 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,2), 'M/s l2res', round(d['value_l2_resident']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'cpu', round(d['cpu_baseline']['value']/1e6,3), d['cpu_baseline']['sample'], 'roof', round(d['roofline']['frac'],4), {k:round(v,4) for k,v in d['phase_ms'].items()})"
done
timeout 600 python bench.py --workload batched_pyramids --steps 30 --warmup 20 --no-cpu 2>&1 | tail -1 > gpurun_out/table_batched.json
python -c "
import json
d=json.load(open('gpurun_out/table_batched.json'))
print('batched', 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'roof', round(d['roofline']['frac'],4), d['config'], {k:round(v,4) for k,v in d['phase_ms'].items()})"
