set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/t_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "suite rc=$?" >> gpurun_out/t_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
timeout 900 python bench.py --all-configs > gpurun_out/bench_all.json 2> gpurun_out/bench_all.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_final_launches.csv python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:go2_step_kernel -s 10 -c 1 -o gpurun_out/r02_step_kernel -f python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/t_noalloc_alone.log gpurun_out/t_gpu.log gpurun_out/smoke.log; cat gpurun_out/bench_default.json
