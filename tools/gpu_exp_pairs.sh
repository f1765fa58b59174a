set -x
mkdir -p gpurun_out
python tools/exp_pairs.py > gpurun_out/exp_pairs.log 2>&1; cat gpurun_out/exp_pairs.log
python tools/exp_trace.py > gpurun_out/trace.log 2>&1; tail -n 70 gpurun_out/trace.log
