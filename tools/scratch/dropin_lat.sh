#!/bin/bash
# latency of the C++ drop-in classes on one synthetic 720p stereo pair
cd "$GRAFT_REPO_ROOT"
python - <<'PY'
from lpslam_b200.synth import synth_stereo_pair
l, r = synth_stereo_pair(720, 1280, 4242)
l.tofile("/tmp/l.raw"); r.tofile("/tmp/r.raw")
PY
for v in 0 1; do LPSLAM_CUDA_LAZY_PYRAMID=$v cpp/build/dropin_smoke 1280 720 2000 /tmp/l.raw /tmp/r.raw latency:200 84.0 0.12; done
