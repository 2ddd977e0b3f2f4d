cd /root/repo
# weak scaling of the headline workload (one many_pyramids world per rank) and strong scaling of the batched config (8192 worlds over N ranks)
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 100 --warmup 30 --no-cpu 2>&1 | grep '^{' | tail -1 > gpurun_out/scale_many_$n.json
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --workload batched_pyramids --steps 60 --warmup 30 --no-cpu 2>&1 | grep '^{' | tail -1 > gpurun_out/scale_batched_$n.json
done
python - <<'PY'
import json
for wl in ("many","batched"):
    for n in (2,4,8):
        try:
            d=json.load(open(f'gpurun_out/scale_{wl}_{n}.json'))
            print(wl, n, 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,1), 'e2e', round(d['e2e']['value']/1e6,1), d.get('scaling'), d['config'].get('workload'), d['config'].get('worlds_per_gpu'), d['clocks'])
        except Exception as e:
            print(wl, n, 'FAILED', e)
PY
