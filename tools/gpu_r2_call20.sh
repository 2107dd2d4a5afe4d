# round 2, call 20 (one pipeline for frames of at most two launches): full GPU suite (no-alloc test collected last), that test alone, probes, per-frame counters,
# bench lines, launch list of the bench command, one full capture of the headline frame and of the 10 M frame
T="timeout -k 10"
$T 900 python -m pytest tests -m gpu -q -x --timeout 400 --durations=8 > gpurun_out/r2c20_pytest.log 2>&1; tail -14 gpurun_out/r2c20_pytest.log
$T 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 150 > gpurun_out/r2c20_noalloc_alone.log 2>&1; tail -2 gpurun_out/r2c20_noalloc_alone.log
rm -f gpurun_out/r2c20_probe.jsonl
for sc in "synth2236 8 1920 1080 4" "synth300 16 1920 1080 4" "indoor_tex 64 1920 1080 3" "indoor_tex 16 3840 2160 8"; do
  $T 150 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c20_probe.jsonl
done
cat gpurun_out/r2c20_probe.jsonl
M=smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
cnt() { # scene w h b
  $T 300 ncu --metrics $M --clock-control none --print-units base --csv --log-file gpurun_out/r2_ncu_$1_$2x$3_b$4.csv \
     python tools/frames_only.py $1 3 $2 $3 $4 > gpurun_out/ncu_cnt.log 2>&1
  python tools/ncu_counters.py gpurun_out/r2_ncu_$1_$2x$3_b$4.csv $1 3 $2 $3 $4 gpurun_out/r2_counters.json > /dev/null
}
rm -f gpurun_out/r2_counters.json
cnt indoor_tex 1920 1080 3; cnt indoor_tex 1920 1080 4; cnt indoor_tex 640 360 4; cnt indoor_tex 3840 2160 8; cnt synth2236 1920 1080 4
mkdir -p profiles; cp gpurun_out/r2_counters.json profiles/r2_counters.json
$T 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c20_bench_n1.json 2> gpurun_out/r2c20_bench_n1.err; tail -2 gpurun_out/r2c20_bench_n1.err
$T 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2c20_bench_ref.json 2> gpurun_out/r2c20_bench_ref.err; tail -2 gpurun_out/r2c20_bench_ref.err
$T 400 python bench.py --workload cfg3 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r2c20_bench_cfg3_10M.json 2> gpurun_out/r2c20_bench_cfg3.err; tail -2 gpurun_out/r2c20_bench_cfg3.err
$T 400 python bench.py --workload cfg2 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r2c20_bench_cfg2_4k_n1.json 2> gpurun_out/r2c20_bench_cfg2.err; tail -2 gpurun_out/r2c20_bench_cfg2.err
$T 300 python bench.py --workload cfg0 --steps 16 --warmup 3 --no-cpu-baseline --sustain-s 2 > gpurun_out/r2c20_bench_cfg0.json 2> gpurun_out/r2c20_bench_cfg0.err
$T 300 python bench.py --workload cfg1b4 --steps 16 --warmup 3 --no-cpu-baseline --sustain-s 2 > gpurun_out/r2c20_bench_cfg1b4.json 2> gpurun_out/r2c20_bench_cfg1b4.err
for f in gpurun_out/r2c20_bench_*.json; do echo $f; cut -c1-400 $f; done
# launch list of the bench command (per-launch durations; cold-cache, serialised)
$T 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2c20_bench_launches.csv \
   python bench.py --steps 4 --warmup 3 --no-cpu-baseline --sustain-s 0 > gpurun_out/ncu_ll.log 2>&1; tail -1 gpurun_out/ncu_ll.log | cut -c1-200
# full captures: one headline frame (6 launches), one 10 M-triangle frame (14 launches)
$T 500 ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 4 --launch-count 2 \
   -o gpurun_out/prof_r2_frame -f python tools/frames_only.py indoor_tex 3 1920 1080 3 > gpurun_out/ncu_full.log 2>&1; tail -1 gpurun_out/ncu_full.log | cut -c1-200
$T 600 ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 14 --launch-count 14 \
ls -la gpurun_out | head -50
