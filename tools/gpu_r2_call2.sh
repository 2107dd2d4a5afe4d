# round 2, call 2: lane statistics + ncu full captures (k_extend on 10 M triangles / 180 k triangles, headline k_path)
set -x
python -m pytest tests -m gpu -q --timeout 1800 -x > gpurun_out/r2c2_pytest.log 2>&1; tail -3 gpurun_out/r2c2_pytest.log
for sc in synth2236 synth300; do
  python tools/probe.py $sc 8 1920 1080 4 2>&1 | tail -1 >> gpurun_out/r2c2_probe.jsonl
done
python tools/probe.py indoor_tex 32 1920 1080 3 2>&1 | tail -1 >> gpurun_out/r2c2_probe.jsonl
cat gpurun_out/r2c2_probe.jsonl
ncu --set full --clock-control none --import-source on -k regex:k_extend --launch-skip 6 --launch-count 2 \
    -o gpurun_out/r2c2_extend_10M -f python tools/probe.py synth2236 2 1920 1080 4 > gpurun_out/ncu1.log 2>&1; tail -2 gpurun_out/ncu1.log | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:k_extend --launch-skip 6 --launch-count 2 \
    -o gpurun_out/r2c2_extend_180k -f python tools/probe.py synth300 2 1920 1080 4 > gpurun_out/ncu2.log 2>&1; tail -2 gpurun_out/ncu2.log | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:k_path --launch-skip 18 --launch-count 6 \
    -o gpurun_out/r2c2_headline -f python tools/probe.py indoor_tex 2 1920 1080 3 > gpurun_out/ncu3.log 2>&1; tail -2 gpurun_out/ncu3.log | cut -c1-200
ls -la gpurun_out/*.ncu-rep
