PTC_LIBRARY=/root/repo/cuda-pathtracer_b200/_variants/f32.so python -m pytest tests -m gpu -q -x --timeout 900 -k "not oracle_bvh and not scene_info" 2>&1 | tail -2
for rep in 1 2; do
for v in "" f32; do
  lib=""; [ -n "$v" ] && lib=/root/repo/cuda-pathtracer_b200/_variants/$v.so
  echo "== ${v:-default}"
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
done
done
