# A/B of the round's kernel changes: camera-basis hoist, deferred epilogue, FFMA2 node test
python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
V=/root/repo/cuda-pathtracer_b200/_variants
run() { # label lib defer
  echo "== $1"
  PTC_LIBRARY=$2 PTC_DEFER=$3 python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_LIBRARY=$2 PTC_DEFER=$3 python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
  PTC_LIBRARY=$2 PTC_DEFER=$3 python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
  PTC_LIBRARY=$2 PTC_DEFER=$3 python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
}
run "baseline (no hoist, epilogue in place)" $V/nohoist.so 0
run "hoist" "" 0
run "hoist + deferred epilogue" "" 1
run "hoist + deferred epilogue + FFMA2" $V/f2.so 1
for rep in 2 3; do
for v in "nohoist 0" "default 0" "default 1" "f2 1"; do
  set -- $v; lib=$V/$1.so; [ $1 = default ] && lib=""
  echo "== rep $rep: $1 defer=$2"
  PTC_LIBRARY=$lib PTC_DEFER=$2 python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
done
done
