#!/bin/bash
# quadtree phase clocks (profile build) and the stage time with 512 / 1024 threads per block; restores the normal build
cd "$GRAFT_REPO_ROOT"
LPSLAM_NVCC_FLAGS="-DLPS_QT_PROFILE" python -m lpslam_b200.build --force > /dev/null 2>&1 || echo build failed
python tools/qt_phases.py 2>&1 | tail -12
for t in 512 1024; do
  LPSLAM_NVCC_FLAGS="-DLPS_QT_THREADS=$t" python -m lpslam_b200.build --force > /dev/null 2>&1 || echo build failed
  echo "threads $t: $(python tools/stage_times.py 32 30 2>&1 | grep -o "quadtree.: [0-9.]*") pair: $(python tools/scratch/pair_latency.py 2>&1 | tail -1)"
done
