mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_pre.json 2> gpurun_out/ncu_pre.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 260 -c 200 --csv --log-file gpurun_out/r02_final_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:go2_step_kernel -s 300 -c 1 -o gpurun_out/r02_step_kernel -f python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_full.log | cut -c1-200
