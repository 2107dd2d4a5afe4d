# round 2, call 4: all GPU tests (incl. group) + node layout A/B (4 loads of 32 bytes vs 7 loads per node)
python -m pytest tests -m gpu -q --timeout 1800 > gpurun_out/r2c4_pytest.log 2>&1; tail -4 gpurun_out/r2c4_pytest.log
V=/root/repo/cuda-pathtracer_b200/_variants
for v in default lay0 tbpf2 lay0tbpf2 tb; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py synth2236 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c4_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py synth300 16 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c4_probe.jsonl
done
for v in default lay0 pathtb default lay0; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py indoor_tex 64 1920 1080 3 2>&1 | tail -1 >> gpurun_out/r2c4_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py indoor_tex 16 3840 2160 8 2>&1 | tail -1 >> gpurun_out/r2c4_probe.jsonl
done
cat gpurun_out/r2c4_probe.jsonl
