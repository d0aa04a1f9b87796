D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  bash profiles/tools/ab.sh fast448 GO2_B200_LIB=$D/libgo2_b200_fast.so
  bash profiles/tools/ab.sh fast224x2 GO2_B200_LIB=$D/variants/libgo2_224.so
done
