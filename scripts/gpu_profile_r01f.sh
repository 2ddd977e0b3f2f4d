cd /root/repo
# plain run first (must exit 0), then the launch list and one --set full capture of the dominant kernel, then the cluster solver on joint_grid
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/r01f_plain.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 300 --csv --log-file gpurun_out/r01f_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/r01f_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_solver_islands -s 5 -c 2 -o gpurun_out/prof_islands_r01f python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/r01f_ncu2.log 2>&1
timeout 300 python bench.py --workload joint_grid --steps 4 --warmup 3 --no-cpu > gpurun_out/r01f_plain_jg.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_solver_cluster -s 3 -c 1 -o gpurun_out/prof_cluster_r01f python bench.py --workload joint_grid --steps 4 --warmup 3 --no-cpu > gpurun_out/r01f_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep
