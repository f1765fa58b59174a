set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decide_kernel -s ${SKIP:-20} -c 1 -o gpurun_out/${OUT:-decide_lo} -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_decide.log 2>&1; echo "ncu full rc=$?"
