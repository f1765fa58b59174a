mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t.log; tail -n 6 gpurun_out/t.log
{
python tools/parity_full.py 1 1500
python tools/parity_full.py 0 300
python tools/parity_full.py 2 200
python tools/parity_full.py 3 16
} > gpurun_out/parity_final.json 2> gpurun_out/parity_final.err
cat gpurun_out/parity_final.json; tail -n 3 gpurun_out/parity_final.err
python bench.py --steps 3 --warmup 3 > gpurun_out/b.json 2> gpurun_out/b.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/b.json
