python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1
grep -n "^E  \|^____\|passed\|failed" gpurun_out/pytest_gpu.log | grep -v "where\|+ " | head -20
nproc
PTC_BUILD_TIMING=1 python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep -v Warn
python tools/frame_probe.py indoor_tex 64 2>&1 | grep -v Warn
