# round-2 profile artefacts: launch list of one bench run + one --set full capture of the dominant kernel (lo_kernel)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/r2_bench_profile_plain.json 2> gpurun_out/r2_bench_profile_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 350 -c 700 --csv --log-file gpurun_out/r2_launches_whole_path.csv \
    python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lo_kernel -s 20 -c 1 -o gpurun_out/r2_lo_kernel -f \
    python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
