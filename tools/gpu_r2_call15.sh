# round 2, call 15: iterations per stage by frame size (first-stage depth)
T="timeout -k 10"
for sc in indoor_tex color_sample_plain island_plain synth300 synth2236; do $T 300 python tools/sweep_first.py $sc 2>&1 | grep '^{' | tee -a gpurun_out/r2_sweep_first.jsonl; done
