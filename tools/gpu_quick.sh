python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1
grep -n "^E  \|^____\|passed\|failed" gpurun_out/pytest_gpu.log | grep -v "where\|+ " | head -20
for d in 0 1 2 3; do
echo "== PTC_DUAL=$d"
PTC_DUAL=$d python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
PTC_DUAL=$d python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
PTC_DUAL=$d python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
PTC_DUAL=$d python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
PTC_DUAL=$d python tools/frame_probe.py indoor_tex 64 640 360 4 2>&1 | grep ms/frame
done
