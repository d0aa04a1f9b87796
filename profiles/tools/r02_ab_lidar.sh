# interleaved A/B in ONE call on ONE box: the product library vs the same source built with -DG2_NO_LIDAR
V=go2-sim2real-locomotion-rl_b200/go2_b200/variants/libgo2_nolidar.so
for i in 1 2 3; do
  bash profiles/tools/ab.sh product
  bash profiles/tools/ab.sh nolidar GO2_B200_LIB=$V
done
bash profiles/tools/ab2.sh product stair 8192
bash profiles/tools/ab2.sh nolidar stair 8192 GO2_B200_LIB=$V
bash profiles/tools/ab2.sh product walk 4096
bash profiles/tools/ab2.sh nolidar walk 4096 GO2_B200_LIB=$V
