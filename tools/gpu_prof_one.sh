# usage: KERNEL=regex SKIP=n OUT=name bash tools/gpu_prof_one.sh   (one --set full capture of one launch)
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${KERNEL} -s ${SKIP:-0} -c 1 -o gpurun_out/${OUT} -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_${OUT}.log 2>&1; echo "rc=$?"
ls -la gpurun_out
