mkdir -p gpurun_out
cp madpose_b200/libmadpose_b200.so /tmp/lib_c12.so
{
for v in c12 c8 c6 c12 c8 c6; do
  if [ $v = c12 ]; then cp /tmp/lib_c12.so madpose_b200/libmadpose_b200.so; else cp build_exp/libmadpose_b200_$v.so madpose_b200/libmadpose_b200.so; fi
  echo "variant $v"; python tools/exp_groups.py 2 2>&1 | tail -n 1
  python tools/exp_pairs.py 188 2>&1 | tail -n 1
done
} > gpurun_out/exp_minctas.log 2>&1
cp /tmp/lib_c12.so madpose_b200/libmadpose_b200.so
cat gpurun_out/exp_minctas.log
