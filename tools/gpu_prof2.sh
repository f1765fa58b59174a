set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lo_kernel -s 5 -c 1 -o gpurun_out/lo_v3 -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_lo.log 2>&1; echo "rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:minimal_batch -s 0 -c 1 -o gpurun_out/minimal_v3 -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_min.log 2>&1; echo "rc=$?"
