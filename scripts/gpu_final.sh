cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench.py --impl reference > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err; tail -c 600 gpurun_out/final_reference.json
timeout 600 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -c 1500 gpurun_out/final_bench.json
timeout 300 python scripts/gpu_mutator_cost.py 2>&1 | tail -12
