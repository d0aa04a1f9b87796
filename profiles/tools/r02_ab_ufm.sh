D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2 3; do
  bash profiles/tools/ab.sh fast GO2_B200_LIB=$D/libgo2_b200_fast.so
  bash profiles/tools/ab.sh use_fast_math GO2_B200_LIB=$D/variants/libgo2_ufm.so
done
cp $D/variants/libgo2_ufm.so $D/libgo2_b200.so
python profiles/tools/parity_report.py 2>&1 | tail -12
