# e2e leg (go2_step_host, host buffers): action upload on its own stream beside the kernel's opening phase vs on the caller's stream before the kernel
timeout 600 python -m pytest tests -x -q -m gpu -k "host or e2e or abi" 2>&1 | tail -2
for i in 1 2 3; do
  for v in overlap serial; do
    if [ $v = serial ]; then export GO2_B200_SERIAL_UPLOAD=1; else unset GO2_B200_SERIAL_UPLOAD; fi
    timeout 300 python bench.py --steps 30 --warmup 30 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$v', 'device', round(d['value']/1e6,2), 'M  e2e', round(d['e2e']['value']/1e6,2), 'M  e2e ms', round(d['e2e']['ms_per_step'],4))"
  done
done
