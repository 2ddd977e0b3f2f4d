cd /root/repo
B2CUDA_DEBUG_CHECKSUMS=2 timeout 200 python scripts/gpu_pairs_stats.py
for wl in "$@"; do timeout 300 python bench.py --workload $wl --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 > /tmp/o.json; python -c "
import json
d=json.load(open('/tmp/o.json')); print(d['config']['workload'], round(d['ms_per_step'],4), round(d['value']/1e6,1), round(d['e2e']['value']/1e6,1), {k:round(v,4) for k,v in d['phase_ms'].items()})"; done
