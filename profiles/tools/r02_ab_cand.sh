# narrow phase on "the k-th candidate of every lane" (two-pass sweep) vs the narrow phase inside the pair sweep (variants/libgo2_base.so)
D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  bash profiles/tools/ab.sh base GO2_B200_LIB=$D/variants/libgo2_base.so
  bash profiles/tools/ab.sh cand GO2_B200_LIB=$D/libgo2_b200_fast.so
done
for w in stair crouch jump; do
python bench.py --workload $w --steps 30 --warmup 30 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-120
GO2_B200_LIB=$D/variants/libgo2_base.so python bench.py --workload $w --steps 30 --warmup 30 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-120
done
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
