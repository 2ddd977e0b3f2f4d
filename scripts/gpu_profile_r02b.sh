cd /root/repo
# round 2, build b (event-free step, graph launch, bullets, overflow joints): plain runs first (must exit 0), then the launch list, then --set full captures of the island solver and the batch kernel
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02b_bench_n1.json 2> gpurun_out/r02b_bench_n1.err || exit 1
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02b_bench_ref.json 2>&1
for wl in large_pyramid tumbler joint_grid; do timeout 600 python bench.py --workload $wl --steps 200 --warmup 60 --no-batched > gpurun_out/r02b_table_$wl.json 2> gpurun_out/r02b_table_$wl.err; done
timeout 600 python bench.py --steps 200 --warmup 60 > gpurun_out/r02b_table_many_pyramids.json 2> gpurun_out/r02b_table_many_pyramids.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 400 --csv --log-file gpurun_out/r02b_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu --worlds 1024 > gpurun_out/r02b_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solver_islands -s 5 -c 2 -o gpurun_out/prof_islands_r02b python bench.py --steps 4 --warmup 3 --no-cpu --no-batched > gpurun_out/r02b_ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_batch_ -s 9 -c 3 -o gpurun_out/prof_batch_r02b python scripts/batch_bench.py --worlds 8192 --steps 4 --warmup 3 > gpurun_out/r02b_ncu3.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step_front -s 5 -c 1 -o gpurun_out/prof_front_r02b python bench.py --steps 4 --warmup 3 --no-cpu --no-batched > gpurun_out/r02b_ncu4.log 2>&1
# gpurun brings back at most 64 MiB: the raw pages go home as CSV, only the island solver's report travels
for n in islands batch front; do ncu -i gpurun_out/prof_${n}_r02b.ncu-rep --page raw --csv > gpurun_out/prof_${n}_r02b_raw.csv 2>/dev/null; done
rm -f gpurun_out/prof_batch_r02b.ncu-rep gpurun_out/prof_front_r02b.ncu-rep
ls -la gpurun_out/ | tail -20
