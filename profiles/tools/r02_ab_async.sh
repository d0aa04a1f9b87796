# warps free of each other (-DG2_WARP_ASYNC: no phase barriers, warp-level Newton loop vote) vs the phase-locked product, and 3 x 160-thread blocks
D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  bash profiles/tools/ab.sh fast224x2 GO2_B200_LIB=$D/libgo2_b200_fast.so
  for v in async224 async448 async128 sync160x3; do bash profiles/tools/ab.sh $v GO2_B200_LIB=$D/variants/libgo2_$v.so; done
done
