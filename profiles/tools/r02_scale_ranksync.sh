N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --rank-sync --no-cpu-baseline > gpurun_out/r02_scale_ranksync_${N}gpu.json 2> gpurun_out/r02_scale_ranksync_${N}gpu.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02_scale_ranksync_${N}gpu.json").read().strip().splitlines()[-1])
print("ranksync", d["n_gpus"], round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3))
PY
