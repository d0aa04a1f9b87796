D=go2-sim2real-locomotion-rl_b200/go2_b200
cp $D/libgo2_b200.so /tmp/product.so
for v in fastdivonly fastsqrtonly fastdiv; do
  bash profiles/tools/ab.sh product
  bash profiles/tools/ab.sh $v GO2_B200_LIB=$D/variants/libgo2_$v.so
done
for v in product fastdivonly fastsqrtonly fastdiv; do
  echo "== parity of $v"
  if [ $v = product ]; then cp /tmp/product.so $D/libgo2_b200.so; else cp $D/variants/libgo2_$v.so $D/libgo2_b200.so; fi
  python profiles/tools/parity_report.py 2>&1 | tail -12
done
