mkdir -p gpurun_out
{
for b in 64 128 256; do MP_BATCH=$b python tools/exp_groups.py 2 2>&1 | tail -n 1; done
MP_B_AFTER_LO=32 MP_BATCH=256 python tools/exp_groups.py 2 2>&1 | tail -n 1
MP_B_AFTER_LO=128 MP_BATCH=256 python tools/exp_groups.py 2 2>&1 | tail -n 1
python tools/exp_pairs.py 188 1500 6000
} > gpurun_out/exp_batch.log 2>&1
cat gpurun_out/exp_batch.log
