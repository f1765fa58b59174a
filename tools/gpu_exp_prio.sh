mkdir -p gpurun_out
{
for m in 0 1 2 0 1 2; do echo "MP_PRIO=$m"; MP_PRIO=$m python tools/exp_groups.py 2 3 2>&1 | tail -n 2; done
} > gpurun_out/exp_prio.log 2>&1
cat gpurun_out/exp_prio.log
