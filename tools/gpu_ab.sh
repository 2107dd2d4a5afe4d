# A/B template: default build against variants built by tools/build_variants.py
V=/root/repo/cuda-pathtracer_b200/_variants
VARIANT=${VARIANT:-sign}
PTC_LIBRARY=$V/$VARIANT.so python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -2
for v in default $VARIANT default $VARIANT; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  echo "== $v"
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
  PTC_LIBRARY=$lib python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
done
