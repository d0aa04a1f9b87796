D=go2-sim2real-locomotion-rl_b200/go2_b200
for i in 1 2; do
  bash profiles/tools/ab.sh fastdiv GO2_B200_LIB=$D/variants/libgo2_fastdiv.so
  bash profiles/tools/ab.sh fastftz GO2_B200_LIB=$D/variants/libgo2_fastftz.so
done
cp $D/variants/libgo2_fastftz.so $D/libgo2_b200.so
python profiles/tools/parity_report.py 2>&1 | tail -12
