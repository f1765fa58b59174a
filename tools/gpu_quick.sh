set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t.log; tail -n 15 gpurun_out/t.log
python bench.py --steps 3 --warmup 3 > gpurun_out/b.json 2> gpurun_out/b.err; echo "bench rc=$?"; cat gpurun_out/b.json; tail -n 5 gpurun_out/b.err
