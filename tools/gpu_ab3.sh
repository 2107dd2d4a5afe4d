# A/B of the shifted schedule (k_wave) against the fused one
python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
run() { # label wave
  echo "== $1"
  PTC_WAVE=$2 python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_WAVE=$2 python tools/frame_probe.py indoor_tex 64 2>&1 | grep ms/frame
  PTC_WAVE=$2 python tools/frame_probe.py indoor_tex 64 1920 1080 4 2>&1 | grep ms/frame
  PTC_WAVE=$2 python tools/frame_probe.py indoor_tex 16 3840 2160 8 2>&1 | grep ms/frame
  PTC_WAVE=$2 python tools/frame_probe.py color_sample_plain 16 3840 2160 3 2>&1 | grep ms/frame
  PTC_WAVE=$2 python tools/frame_probe.py indoor_tex 64 640 360 4 2>&1 | grep ms/frame
}
run "fused (k_path)" 0
run "shifted (k_wave)" 1
python tools/frame_probe.py synth2236 32 1920 1080 4 2>&1 | grep ms/frame
