set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t1.log; tail -3 gpurun_out/t1.log
python bench.py --steps 3 --warmup 3 > gpurun_out/b1.json 2> gpurun_out/b1.err; echo "bench rc=$?"; cat gpurun_out/b1.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/b1_ref.json 2>&1; cat gpurun_out/b1_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches1.csv python bench.py --steps 1 --warmup 1 --pairs 1500 > gpurun_out/ncu1.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decide_kernel -s 1 -c 1 -o gpurun_out/decide_v1 -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu2.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:minimal_batch -s 0 -c 1 -o gpurun_out/minimal_v1 -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu3.log 2>&1; echo "ncu full rc=$?"
