for lib in "$@"; do
  echo "== $lib"
  MADPOSE_B200_LIB=$PWD/madpose_b200/$lib python tools/exp_warp_from.py 0
done
