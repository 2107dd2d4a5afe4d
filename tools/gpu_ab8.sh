# A/B: packed pixel + env reuse (default) vs previous commit; path kernels at 3 CTAs/SM and/or packed FMA
V=/root/repo/cuda-pathtracer_b200/_variants
python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -2
for v in prev default p3 p3f2 f2 prev default; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  echo "== $v"
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
done
for v in prev default p3f2; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  echo "== $v"
  PTC_LIBRARY=$lib python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
done
