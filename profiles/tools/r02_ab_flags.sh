D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2; do
  bash profiles/tools/ab.sh fast GO2_B200_LIB=$D/libgo2_b200_fast.so
  for v in vec exp restrict; do bash profiles/tools/ab.sh $v GO2_B200_LIB=$D/variants/libgo2_$v.so; done
done
