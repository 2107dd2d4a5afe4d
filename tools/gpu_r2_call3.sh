# round 2, call 3: memory-wait restructuring of the traversal step (batched triangle loads, early node header)
V=/root/repo/cuda-pathtracer_b200/_variants
for v in default pf3 tb tbpf3 tbpf2 tbpf3ext4 ext2 tbpf3ext2; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py synth2236 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c3_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py synth300 16 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c3_probe.jsonl
done
for v in default pathtb pathpf3 pathtbpf3 default pathtbpf3; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py indoor_tex 64 1920 1080 3 2>&1 | tail -1 >> gpurun_out/r2c3_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib python tools/probe.py indoor_tex 16 3840 2160 8 2>&1 | tail -1 >> gpurun_out/r2c3_probe.jsonl
done
cat gpurun_out/r2c3_probe.jsonl
