# terrain narrow phase: feet (spheres) run MPR only against prisms within one radius in 3-D (exact point-prism distance) vs within one radius of the footprint in xy
D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  for w in stair stair_info; do
    for v in cull3d cull2d; do
      if [ $v = cull2d ]; then export GO2_B200_LIB=$D/variants/libgo2_cull2d.so; else unset GO2_B200_LIB; fi
      timeout 300 python bench.py --workload $w --steps 30 --warmup 30 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$w $v', round(d['value']/1e6,3), 'M  kernel_ms', round(d['roofline']['kernel_ms_avg'],3), ' e2e', round(d['e2e']['value']/1e6,3))"
    done
  done
done
unset GO2_B200_LIB
timeout 600 python -m pytest tests -x -q -m gpu -k "stair or terrain or golden or shipped" 2>&1 | tail -2
