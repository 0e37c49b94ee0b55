#!/bin/bash
# Builds and runs scripts/cpp/bench_operator.cpp at the three camera shapes (run on the GPU box from the repo root).
set -e
g++ -O2 -std=c++14 -w -Iinclude -Ioracle/shim scripts/cpp/bench_operator.cpp -Lorbslam2_b200 -lorbx -Wl,-rpath,$PWD/orbslam2_b200 -o /tmp/bench_operator
python - <<'PY'
import sys; sys.path.insert(0, '.')
from orbslam2_b200.synth import synth_image
for w, h in ((640, 480), (752, 480), (1241, 376)):
    synth_image(w, h, 0).tofile(f"/tmp/img_{w}x{h}.raw")
PY
/tmp/bench_operator 640 480 1000 /tmp/img_640x480.raw
/tmp/bench_operator 752 480 1200 /tmp/img_752x480.raw
/tmp/bench_operator 1241 376 2000 /tmp/img_1241x376.raw
