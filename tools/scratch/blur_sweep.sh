#!/bin/bash
# on the GPU box: rebuild blur_tma.cu with different block widths / register caps and time the blur stage (median of 20 batches)
cd "$GRAFT_REPO_ROOT"
for cfg in "1 6" "1 8" "2 6" "2 7"; do
  set -- $cfg
  LPSLAM_NVCC_FLAGS="-DLPS_BT_WARPS=$1 -DLPS_BT_MINB=$2" python -m lpslam_b200.build --force > /dev/null 2>&1 || echo build failed
  for s in 40 50 60 80; do
    echo "warps $1 minb $2 seg $s: $(LPS_BLUR_SEG_ROWS=$s python tools/stage_times.py 32 30 2>&1 | grep -o "blur.: [0-9.]*")"
  done
done
