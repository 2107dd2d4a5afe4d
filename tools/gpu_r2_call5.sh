# round 2, call 5: GPU tests (every command under its own timeout), node layout A/B, ncu counters per workload, bench lines
T="timeout -k 10"
$T 600 python -m pytest tests -m gpu -q -x --timeout 300 > gpurun_out/r2c5_pytest.log 2>&1; tail -4 gpurun_out/r2c5_pytest.log
$T 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 100 > gpurun_out/r2c5_noalloc_alone.log 2>&1; tail -2 gpurun_out/r2c5_noalloc_alone.log
V=/root/repo/cuda-pathtracer_b200/_variants
for v in default lay0 tbpf2 lay0tbpf2 tb; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib $T 120 python tools/probe.py synth2236 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c5_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib $T 120 python tools/probe.py synth300 16 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c5_probe.jsonl
done
for v in default lay0 pathtb default lay0; do
  lib=$V/$v.so; [ $v = default ] && lib=""
  PTC_TAG=$v PTC_LIBRARY=$lib $T 120 python tools/probe.py indoor_tex 64 1920 1080 3 2>&1 | tail -1 >> gpurun_out/r2c5_probe.jsonl
  PTC_TAG=$v PTC_LIBRARY=$lib $T 120 python tools/probe.py indoor_tex 16 3840 2160 8 2>&1 | tail -1 >> gpurun_out/r2c5_probe.jsonl
done
cat gpurun_out/r2c5_probe.jsonl
M=smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
cnt() { # scene w h b
  $T 300 ncu --metrics $M --clock-control none --print-units base --csv --log-file gpurun_out/r2_ncu_$1_$2x$3_b$4.csv \
     python tools/frames_only.py $1 3 $2 $3 $4 > gpurun_out/ncu_cnt.log 2>&1
  python tools/ncu_counters.py gpurun_out/r2_ncu_$1_$2x$3_b$4.csv $1 3 $2 $3 $4 gpurun_out/r2_counters.json
}
cnt indoor_tex 1920 1080 3; cnt indoor_tex 1920 1080 4; cnt indoor_tex 640 360 4; cnt indoor_tex 3840 2160 8; cnt synth2236 1920 1080 4
cp gpurun_out/r2_counters.json profiles/r2_counters.json
$T 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c5_bench_n1.json 2> gpurun_out/r2c5_bench_n1.err; tail -2 gpurun_out/r2c5_bench_n1.err; cut -c1-1500 gpurun_out/r2c5_bench_n1.json
