cd /root/repo
for worlds in 512 768 1024 1536; do for mode in cta warp; do
B2CUDA_ISLAND_MODE=$mode timeout 300 python bench.py --workload batched_pyramids --worlds $worlds --steps 60 --warmup 30 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$worlds $mode', round(d['ms_per_step'],4), round(d['ms_per_step_l2_resident'],4), round(d['phase_ms']['solver_kernel'],4))"; done; done
