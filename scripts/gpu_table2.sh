cd /root/repo
wl=tumbler
timeout 900 python bench.py --workload $wl --steps 200 --warmup 60 --cpu-steps 20 2>&1 | tail -1 > gpurun_out/table_$wl.json
python -c "
import json,sys
d=json.load(open('gpurun_out/table_$wl.json'))
print(d['config']['workload'], 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,2), 'M/s l2res', round(d['value_l2_resident']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'cpu', round(d['cpu_baseline']['value']/1e6,3), d['cpu_baseline']['sample'], 'roof', round(d['roofline']['frac'],4), {k:round(v,4) for k,v in d['phase_ms'].items()})" || cut -c1-300 gpurun_out/table_$wl.json
