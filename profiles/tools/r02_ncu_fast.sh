mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:go2_step_kernel -s 270 -c 1 -o gpurun_out/r02_step_kernel_fast -f python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_fast.log 2>&1
ls -la gpurun_out/r02_step_kernel_fast.ncu-rep
