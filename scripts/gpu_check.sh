cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for wl in tumbler large_pyramid joint_grid many_pyramids; do timeout 300 python bench.py --workload $wl --steps 32 --warmup 20 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['config']['workload'], d['ms_per_step'], d['value'], d.get('phase_ms'), d.get('e2e'))"; done
