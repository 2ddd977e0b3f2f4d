cd /root/repo
# round 2, final build (r02c: r02b + shared-memory cluster solver, register-resident colouring rounds, scrambled union-find, contact growth):
# plain runs first (must exit 0), then the launch list, then --set full captures; raw pages go home as CSV (gpurun brings back 64 MiB at most)
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02c_bench_n1.json 2> gpurun_out/r02c_bench_n1.err || exit 1
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02c_bench_ref.json 2>&1
for wl in large_pyramid tumbler joint_grid; do timeout 600 python bench.py --workload $wl --steps 200 --warmup 60 --no-batched > gpurun_out/r02c_table_$wl.json 2> gpurun_out/r02c_table_$wl.err; done
timeout 600 python bench.py --steps 200 --warmup 60 > gpurun_out/r02c_table_many_pyramids.json 2> gpurun_out/r02c_table_many_pyramids.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 400 --csv --log-file gpurun_out/r02c_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu --worlds 1024 > gpurun_out/r02c_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solver_islands -s 5 -c 2 -o gpurun_out/prof_islands_r02c python bench.py --steps 4 --warmup 3 --no-cpu --no-batched > gpurun_out/r02c_ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solver_cluster -s 5 -c 1 -o gpurun_out/prof_cluster_r02c python bench.py --workload large_pyramid --steps 4 --warmup 3 --no-cpu --no-batched > gpurun_out/r02c_ncu3.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step_front -s 5 -c 1 -o gpurun_out/prof_front_r02c python bench.py --steps 4 --warmup 3 --no-cpu --no-batched > gpurun_out/r02c_ncu4.log 2>&1
for n in islands cluster front; do ncu -i gpurun_out/prof_${n}_r02c.ncu-rep --page raw --csv > gpurun_out/prof_${n}_r02c_raw.csv 2>/dev/null; done
rm -f gpurun_out/prof_cluster_r02c.ncu-rep gpurun_out/prof_front_r02c.ncu-rep
ls -la gpurun_out/ | tail -24
