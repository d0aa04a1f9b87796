mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu_soa.log 2>&1; echo "suite rc=$?" >> gpurun_out/t_gpu_soa.log
timeout 600 python bench.py --steps 60 --warmup 20 > gpurun_out/bench_soa.json 2> gpurun_out/bench_soa.err
timeout 600 python bench.py --steps 60 --warmup 20 > gpurun_out/bench_soa2.json 2>> gpurun_out/bench_soa.err
timeout 900 python bench.py --all-configs > gpurun_out/bench_soa_all.json 2>> gpurun_out/bench_soa.err
tail -n 3 gpurun_out/t_gpu_soa.log
python - <<'PY'
import json
for f in ("bench_soa.json","bench_soa2.json","bench_soa_all.json"):
    for l in open("gpurun_out/"+f):
        d=json.loads(l); print(f, d["config"]["workload"][:28], d["config"]["envs_per_gpu"], round(d["value"]/1e6,2), "e2e", round(d["e2e"]["value"]/1e6,2), "kernel_ms", round(d["roofline"]["kernel_ms_avg"],4))
PY
