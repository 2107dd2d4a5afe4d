for v in "" ext4 path3 path5; do
  lib=""; [ -n "$v" ] && lib=/root/repo/cuda-pathtracer_b200/_variants/$v.so
  echo "== ${v:-default}"
  PTC_LIBRARY=$lib python tools/sweep2.py indoor_tex 3 1920 1080 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/sweep2.py synth2236 4 1920 1080 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/sweep2.py indoor_tex 8 3840 2160 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/sweep2.py color_sample_plain 3 3840 2160 2>&1 | grep ms/frame
done
