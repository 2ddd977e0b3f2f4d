cd /root/repo
timeout 300 python -m pytest tests -m gpu -x -q -k "move_events or small_pyramid or golden" 2>&1 | tail -2
for rep in 1 2; do for z in 0 1; do
B2CUDA_ZEROCOPY_MOVES=$z timeout 300 python bench.py --steps 200 --warmup 60 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('zerocopy=$z', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']/1e6,1), 'finalize', round(d['phase_ms']['finalize_refit'],4))"; done; done
B2CUDA_ZEROCOPY_MOVES=0 timeout 300 python bench.py --workload batched_pyramids --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('batched zc=0', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']/1e6,1))"
timeout 300 python bench.py --workload batched_pyramids --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('batched zc=1', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']/1e6,1))"
