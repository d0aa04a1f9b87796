mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/f_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/f_gpu.log
timeout 600 python -m pytest tests/test_golden_dr.py -q -m gpu -k fast_math -s 2>&1 | grep "cuda_fast" | cut -c1-260 > gpurun_out/r02_fast_math_parity.txt
timeout 600 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "rc=$?" >> gpurun_out/f_bench.err
timeout 900 python bench.py --all-configs --no-cpu-baseline > gpurun_out/f_bench_all.json 2>> gpurun_out/f_bench.err
bash profiles/tools/r02_ab_ftz.sh > gpurun_out/ab_ftz.txt 2>&1
tail -n 3 gpurun_out/f_gpu.log; cat gpurun_out/r02_fast_math_parity.txt; tail -2 gpurun_out/f_bench.err
python - <<'PY'
import json
for f in ("f_bench.json","f_bench_all.json"):
    for l in open("gpurun_out/"+f):
        d=json.loads(l); print(d["config"]["workload"][:30], d["config"]["envs_per_gpu"], round(d["value"]/1e6,2), "e2e", round(d["e2e"]["value"]/1e6,2), "kernel", round(d["roofline"]["kernel_ms_avg"],3), "precise", round(d.get("precise_math",{}).get("value",0)/1e6,2), "fresh", round(d.get("fresh_after_reset",{}).get("value",0)/1e6,2), "fused", round(d.get("with_policy_fused",{}).get("value",0)/1e6,2), "roll", round(d.get("rollout",{}).get("value",0)/1e6,2))
PY
cat gpurun_out/ab_ftz.txt | tail -12
