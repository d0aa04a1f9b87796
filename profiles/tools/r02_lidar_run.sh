mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu_lidar.log 2>&1; echo "suite rc=$?" >> gpurun_out/t_gpu_lidar.log
timeout 600 python bench.py --workload stair_lidar > gpurun_out/bench_stair_lidar.json 2> gpurun_out/bench_stair_lidar.err
timeout 600 python bench.py --workload stair > gpurun_out/bench_stair_rows.json 2>> gpurun_out/bench_stair_lidar.err
timeout 600 python bench.py > gpurun_out/bench_walk_after.json 2>> gpurun_out/bench_stair_lidar.err
tail -n 4 gpurun_out/t_gpu_lidar.log
python - <<'PY'
import json
for f in ("bench_stair_lidar.json","bench_stair_rows.json","bench_walk_after.json"):
    for l in open("gpurun_out/"+f):
        d=json.loads(l); print(f, d["config"]["workload"][:34], d["config"]["envs_per_gpu"], round(d["value"]/1e6,2), "e2e", round(d["e2e"]["value"]/1e6,2), "kernel_ms", round(d["roofline"]["kernel_ms_avg"],4))
PY
