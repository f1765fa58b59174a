mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/b2.json 2> gpurun_out/b2.err; echo "rc=$?"; grep '^{' gpurun_out/b2.json | cut -c1-300; tail -n 3 gpurun_out/b2.err
