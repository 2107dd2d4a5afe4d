# round 2, call 11 (8 GPUs): group test over 8 devices, headline weak scaling and configs[2] strong
# scaling at N = 8 and N = 4, NCCL-gather form beside it, config 5 in both modes, reference arm
T="timeout -k 10"
nvidia-smi -L | wc -l
$T 600 python -m pytest tests -m gpu -q -x --timeout 400 -k "group or full_image" > gpurun_out/r2c11_pytest.log 2>&1; tail -2 gpurun_out/r2c11_pytest.log
run() { # N port args... ; output name in $OUT
  N=$1; PORT=$2; shift 2
  $T 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/tmp.log 2> gpurun_out/tmp.err
  tail -2 gpurun_out/tmp.err; grep '^{' gpurun_out/tmp.log > gpurun_out/$OUT; cut -c1-260 gpurun_out/$OUT
}
OUT=r2_scale_n8.json run 8 29521 --steps 20 --warmup 5
OUT=r2_scale_n4.json run 4 29522 --steps 20 --warmup 5
OUT=r2_cfg2_n8.json run 8 29523 --workload cfg2 --steps 16 --warmup 3
OUT=r2_cfg2_n4.json run 4 29524 --workload cfg2 --steps 16 --warmup 3
OUT=r2_cfg2_n8_nccl.json run 8 29525 --workload cfg2 --steps 16 --warmup 3 --gather nccl --sustain-s 0
OUT=r2_scale_n8_nccl.json run 8 29526 --steps 20 --warmup 5 --gather nccl --sustain-s 0
$T 300 python bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/r2_bench_ref_n8.json 2>&1; cut -c1-200 gpurun_out/r2_bench_ref_n8.json
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T 400 $TR --master-port 29527 tools/config5.py --mode tiles --frames 8 --out gpurun_out/r2_config5_8k_n8_tiles.json > gpurun_out/cfg5a.log 2>&1; tail -1 gpurun_out/cfg5a.log | cut -c1-300
$T 400 $TR --master-port 29528 tools/config5.py --mode scenes --frames 8 --out gpurun_out/r2_config5_8k_n8_scenes.json > gpurun_out/cfg5b.log 2>&1; tail -1 gpurun_out/cfg5b.log | cut -c1-300
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_scale_n*.json") + glob.glob("gpurun_out/r2_cfg2_n*.json")):
    try:
        d = json.load(open(f))
        print(f, round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d["bit_identity"], d["ms_gather_once"])
    except Exception as e:
        print(f, "unreadable", e)
PY
