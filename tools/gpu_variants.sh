# time the loop with kernel-variant builds of the library (MADPOSE_B200_LIB)
for lib in "$@"; do
  echo "== $lib"
  MADPOSE_B200_LIB=$PWD/madpose_b200/$lib python bench.py --steps 3 --warmup 2 --no-extra 2>/dev/null | python -c "
import json,sys; b=json.loads(sys.stdin.read()); print(round(b['value']), round(b['ms_per_step'],1), round(b['e2e']['value']), {k: round(v,1) for k,v in b['kernel_ms_per_step_single_stream'].items()}, b['rounds_per_step'])"
done
