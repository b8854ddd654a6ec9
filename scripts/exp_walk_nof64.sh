for spec in "" "-DTRC_WALK_EXPERIMENT_NO_F64" "-DTRC_WALK_EXPERIMENT_NO_F64 -DTRC_WALK_MIN_CTAS_FLAT=3 -DTRC_WALK_BATCH=32" "-DTRC_WALK_EXPERIMENT_NO_F64 -DTRC_WALK_MIN_CTAS_FLAT=4 -DTRC_WALK_BATCH=32" "-DTRC_WALK_EXPERIMENT_NO_F64 -DTRC_WALK_MIN_CTAS_FLAT=3 -DTRC_WALK_BATCH=32 -DTRC_WALK_TMA=0"; do
  echo "=== [$spec]"
  TRC_NVCC_EXTRA="$spec" python tracer_b200/csrc/build.py > /dev/null || { echo BUILD FAILED; continue; }
  python scripts/walk_only_probe.py 1e8
done
python tracer_b200/csrc/build.py > /dev/null
