# BVH build cost parameters re-swept (triangle tests got cheaper relative to node tests)
for pc in 0.08 0.12 0.16 0.22 0.30; do
  echo "== PRIM_COST=$pc"
  PTC_PRIM_COST=$pc python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_PRIM_COST=$pc python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
done
for tc in 0.5 2 4; do
  echo "== TRAV_COST=$tc"
  PTC_TRAV_COST=$tc python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_TRAV_COST=$tc python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
done
for pc in 0.12 0.22 0.30; do
  echo "== synth PRIM_COST=$pc"
  PTC_PRIM_COST=$pc python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
done
echo "== synth TRAV_COST=2"
PTC_TRAV_COST=2 python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
