# A/B on the 10 M-triangle scene: L2 prefetch of hit children in the extend kernel, extend kernel at 4 CTAs/SM, refill threshold
V=/root/repo/cuda-pathtracer_b200/_variants
PTC_LIBRARY=$V/pf.so python -m pytest tests -m gpu -q -x --timeout 900 -k "not oracle_bvh and not scene_info" 2>&1 | tail -2
for v in default pf ext4 pfext4 default pf; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  echo "== $v"
  PTC_LIBRARY=$lib python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
done
python tools/sweep_refill.py synth2236 4 1920 1080
