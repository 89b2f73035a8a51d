#!/bin/bash
# on the GPU box: rebuild blur_tma.cu with different block widths and time the blur stage
cd "$GRAFT_REPO_ROOT"
for w in 1 2 4 8; do
  LPSLAM_NVCC_FLAGS="-DLPS_BT_WARPS=$w" python -m lpslam_b200.build --force > /dev/null 2>&1 || echo build failed
  for s in 60 80 100; do
    echo "warps $w seg $s: $(LPS_BLUR_SEG_ROWS=$s python tools/stage_times.py 32 20 2>&1 | grep -o "blur.: [0-9.]*")"
  done
done
