cd /root/repo
cp box2d_b200/csrc/libb2cuda.so /tmp/keep.so
for v in Base Pf Base Pf; do
  cp box2d_b200/csrc/variants/lib$v.so box2d_b200/csrc/libb2cuda.so
  for wl in many_pyramids batched_pyramids; do python bench.py --workload $wl --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$v', d['config']['workload'], round(d['ms_per_step'],4), round(d['ms_per_step_l2_resident'],4), {k:round(v,4) for k,v in d['phase_ms'].items()})"; done
done
cp /tmp/keep.so box2d_b200/csrc/libb2cuda.so
