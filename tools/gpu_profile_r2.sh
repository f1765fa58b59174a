# round-2 profile artefacts: launch list of one bench run + one --set full capture of the dominant kernel (lo_kernel)
# and of the second one (the warp-per-sample 6-point solver); summaries are made here from the exports:
#   python tools/launch_summary.py gpurun_out/r2_launches_whole_path.csv
#   ncu -i gpurun_out/r2_lo_kernel.ncu-rep --page source --csv > lo.csv; nvdisasm -g -c <cubin of mp_kind1.o> > lo.nvd
#   python tools/ncu_lines.py lo.csv lo.nvd lo_kernelILi1
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/r2_bench_profile_plain.json 2> gpurun_out/r2_bench_profile_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 350 -c 900 --csv --log-file gpurun_out/r2_launches_whole_path.csv \
    python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lo_kernel -s 20 -c 1 -o gpurun_out/r2_lo_kernel -f \
    python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ncu --set full --clock-control none --import-source on -k regex:minimal6_warp_batch -s 10 -c 1 -o gpurun_out/r2_w6 -f \
    python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/ncu_w6.log 2>&1
tail -2 gpurun_out/ncu_w6.log
ls -la gpurun_out/*.ncu-rep
