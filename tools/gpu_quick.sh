python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1
grep -n "^E  \|^____\|passed\|failed" gpurun_out/pytest_gpu.log | grep -v "where\|+ " | head -20
for i in 1 2; do
python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
done
