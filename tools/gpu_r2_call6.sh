# round 2, call 6: full GPU suite with the adopted defaults (no-alloc test collected last), then alone; counters; bench
T="timeout -k 10"
$T 600 python -m pytest tests -m gpu -q -x --timeout 300 > gpurun_out/r2c6_pytest.log 2>&1; tail -3 gpurun_out/r2c6_pytest.log
$T 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 100 > gpurun_out/r2c6_noalloc_alone.log 2>&1; tail -2 gpurun_out/r2c6_noalloc_alone.log
for sc in "synth2236 8 1920 1080 4" "synth300 16 1920 1080 4" "indoor_tex 64 1920 1080 3" "indoor_tex 16 3840 2160 8"; do
  $T 120 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c6_probe.jsonl
done
cat gpurun_out/r2c6_probe.jsonl
M=smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
cnt() { # scene w h b
  $T 300 ncu --metrics $M --clock-control none --print-units base --csv --log-file gpurun_out/r2_ncu_$1_$2x$3_b$4.csv \
     python tools/frames_only.py $1 3 $2 $3 $4 > gpurun_out/ncu_cnt.log 2>&1
  python tools/ncu_counters.py gpurun_out/r2_ncu_$1_$2x$3_b$4.csv $1 3 $2 $3 $4 gpurun_out/r2_counters.json > /dev/null
}
rm -f gpurun_out/r2_counters.json
cnt indoor_tex 1920 1080 3; cnt indoor_tex 1920 1080 4; cnt indoor_tex 640 360 4; cnt indoor_tex 3840 2160 8; cnt synth2236 1920 1080 4
cp gpurun_out/r2_counters.json profiles/r2_counters.json
$T 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c6_bench_n1.json 2> gpurun_out/r2c6_bench_n1.err; tail -2 gpurun_out/r2c6_bench_n1.err
$T 400 python bench.py --workload cfg3 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r2c6_bench_cfg3_10M.json 2> gpurun_out/r2c6_bench_cfg3.err; tail -2 gpurun_out/r2c6_bench_cfg3.err
$T 400 python bench.py --workload cfg2 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r2c6_bench_cfg2_4k_n1.json 2> gpurun_out/r2c6_bench_cfg2.err; tail -2 gpurun_out/r2c6_bench_cfg2.err
$T 300 python bench.py --workload cfg0 --steps 16 --warmup 3 --no-cpu-baseline --sustain-s 2 > gpurun_out/r2c6_bench_cfg0.json 2> gpurun_out/r2c6_bench_cfg0.err
$T 300 python bench.py --workload cfg1b4 --steps 16 --warmup 3 --no-cpu-baseline --sustain-s 2 > gpurun_out/r2c6_bench_cfg1b4.json 2> gpurun_out/r2c6_bench_cfg1b4.err
for f in gpurun_out/r2c6_bench_*.json; do echo $f; cut -c1-300 $f; done
