# usage: scripts/ab_probe.sh "<nvcc flags A>" ...   trc_nearest_hit over 1e8 C3 rays for each build variant
for spec in "$@"; do
  echo "=== [$spec]"
  TRC_NVCC_EXTRA="$spec" python tracer_b200/csrc/build.py > /dev/null || { echo BUILD FAILED; continue; }
  python scripts/walk_only_probe.py 1e8 2>&1 | tail -1
done
python tracer_b200/csrc/build.py > /dev/null
