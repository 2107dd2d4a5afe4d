# A/B of experiment builds (tools/build_variants.py) on the two ends of the scene-size range
python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -1
for v in "" tri8 tri12 tri16 tri24; do
  lib=""; [ -n "$v" ] && lib=/root/repo/cuda-pathtracer_b200/_variants/$v.so
  echo "== ${v:-default}"
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
done
PTC_LIBRARY=/root/repo/cuda-pathtracer_b200/_variants/tri12.so python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -1
