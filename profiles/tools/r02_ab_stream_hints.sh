# per-env global rows (state load / store, obs, privileged obs, contact forces) with streaming (evict-first) cache hints vs plain accesses
D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  bash profiles/tools/ab.sh hints GO2_B200_LIB=$D/libgo2_b200_fast.so
  bash profiles/tools/ab.sh plain GO2_B200_LIB=$D/variants/libgo2_nohint.so
done
