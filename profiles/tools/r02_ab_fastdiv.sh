# A/B: IEEE division / sqrt (product) vs -prec-div=false -prec-sqrt=false, interleaved in one call; then the GPU parity suite on the variant
V=go2-sim2real-locomotion-rl_b200/go2_b200/variants/libgo2_fastdiv.so
for i in 1 2 3; do
  bash profiles/tools/ab.sh product
  bash profiles/tools/ab.sh fastdiv GO2_B200_LIB=$V
done
bash profiles/tools/ab2.sh product stair 8192
bash profiles/tools/ab2.sh fastdiv stair 8192 GO2_B200_LIB=$V
cp $V go2-sim2real-locomotion-rl_b200/go2_b200/libgo2_b200.so
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -15
