cd /root/repo
for m in cta warp cta warp; do
  for wl in many_pyramids batched_pyramids; do B2CUDA_ISLAND_MODE=$m python bench.py --workload $wl --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$m', d['config']['workload'], round(d['ms_per_step'],4), round(d['ms_per_step_l2_resident'],4), round(d['e2e']['value']/1e6,1), {k:round(v,4) for k,v in d['phase_ms'].items()})"; done
done
