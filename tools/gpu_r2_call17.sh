# round 2, call 17: full capture of the headline frame on the adopted schedule; pipelines / CTAs per SM / deferred epilogue re-checked on it
T="timeout -k 10"
$T 500 ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 8 --launch-count 4 \
   -o gpurun_out/prof_r2_frame -f python tools/frames_only.py indoor_tex 3 1920 1080 3 > gpurun_out/ncu_full.log 2>&1; tail -1 gpurun_out/ncu_full.log | cut -c1-200
V=/root/repo/cuda-pathtracer_b200/_variants
rm -f gpurun_out/r2c17_probe.jsonl
for sc in "indoor_tex 64 1920 1080 3" "indoor_tex 16 3840 2160 8" "indoor_tex 64 640 360 4"; do
  PTC_TAG=default $T 100 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c17_probe.jsonl
  PTC_TAG=dual0 PTC_DUAL=0 $T 100 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c17_probe.jsonl
  PTC_TAG=dual2 PTC_DUAL=2 $T 100 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c17_probe.jsonl
  PTC_TAG=defer0 PTC_DEFER=0 $T 100 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c17_probe.jsonl
  PTC_TAG=cta3 PTC_LIBRARY=$V/pb3.so $T 100 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c17_probe.jsonl
  PTC_TAG=wave PTC_WAVE=1 $T 100 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c17_probe.jsonl
done
cut -c1-150 gpurun_out/r2c17_probe.jsonl
