# usage: bash profiles/tools/r02_scale_run.sh N   (on an N-GPU box) -> gpurun_out/r02_final_scale_${N}gpu.json (+ reference arm line)
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r02_final_scale_${N}gpu.json 2> gpurun_out/r02_final_scale_${N}gpu.err
echo "rc=$?" >> gpurun_out/r02_final_scale_${N}gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 profiles/tools/pcie_probe.py > gpurun_out/r02_final_pcie_probe_${N}gpu.txt 2>&1
cut -c1-1500 gpurun_out/r02_final_scale_${N}gpu.json
