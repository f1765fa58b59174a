# launch list of one bench run + one --set full capture of the dominant kernel (lo_kernel), after a clean run
set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_clean.json 2> gpurun_out/bench_clean.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lo_kernel -s 9 -c 1 -o gpurun_out/lo_full -f python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_lo_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
