# final build of round 2 (narrow phase over a candidate list, stair-info variant): all-configs bench, launch list, one full ncu capture
mkdir -p gpurun_out
timeout 900 python bench.py --all-configs > gpurun_out/bench_all.json 2> gpurun_out/bench_all.err; echo "rc=$?" >> gpurun_out/bench_all.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_final_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
bash profiles/tools/r02_ncu_fast.sh
tail -2 gpurun_out/bench_all.err; wc -l gpurun_out/r02_final_launches.csv
