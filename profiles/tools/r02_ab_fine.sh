# finer phase lock: block barriers between the sub-phases of a Newton iteration (fineA), plus around env_pre / integrate (fineB)
D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  for v in cand restruct fineA fineB; do bash profiles/tools/ab.sh $v GO2_B200_LIB=$D/variants/libgo2_$v.so; done
done
