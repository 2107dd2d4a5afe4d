# round 2, call 1: new full-size parity tests + k_extend variants on the 10 M-triangle scene
set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python -m pytest tests -m gpu -q --timeout 1800 --durations=8 > gpurun_out/r2c1_pytest.log 2>&1; tail -15 gpurun_out/r2c1_pytest.log
V=/root/repo/cuda-pathtracer_b200/_variants
for v in default pf1 ext4 pf1ext4; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py synth2236 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c1_probe.jsonl
done
for t in 32 256 2048; do
  PTC_TAG=treelet$t PTC_TREELET=$t python tools/probe.py synth2236 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c1_probe.jsonl
  PTC_TAG=pf1_treelet$t PTC_TREELET=$t PTC_LIBRARY=$V/pf1.so python tools/probe.py synth2236 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c1_probe.jsonl
done
for v in default pf1; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py indoor_tex 64 1920 1080 3 2>&1 | tail -1 >> gpurun_out/r2c1_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py synth300 32 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c1_probe.jsonl
done
cat gpurun_out/r2c1_probe.jsonl
