# experiment: two independent single-GPU bench processes at once, unpinned and pinned to disjoint core sets
nproc; lscpu | grep -E "NUMA|Socket|Model name" | head -6; nvidia-smi topo -m 2>/dev/null | head -8
run() { python bench.py --no-extra --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), round(d['ms_per_step'],1), round(d['resident_wall_pairs_per_s']))"; }
echo "-- alone"; CUDA_VISIBLE_DEVICES=0 run alone0
echo "-- two, unpinned"
(CUDA_VISIBLE_DEVICES=0 run u0) & (CUDA_VISIBLE_DEVICES=1 run u1) & wait
N=$(nproc); H=$((N/2))
echo "-- two, pinned 0-$((H-1)) / $H-$((N-1))"
(CUDA_VISIBLE_DEVICES=0 taskset -c 0-$((H-1)) bash -c "$(declare -f run); run p0") & (CUDA_VISIBLE_DEVICES=1 taskset -c $H-$((N-1)) bash -c "$(declare -f run); run p1") & wait
