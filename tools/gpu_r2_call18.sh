# round 2, call 18 (8 GPUs): final multi-GPU lines on the adopted schedule (N = 8, 4, 2 on one box)
T="timeout -k 10"
run() { N=$1; PORT=$2; shift 2
  $T 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/tmp.log 2> gpurun_out/tmp.err
  tail -2 gpurun_out/tmp.err; grep '^{' gpurun_out/tmp.log > gpurun_out/$OUT; cut -c1-160 gpurun_out/$OUT; }
OUT=r2_scale_n8.json run 8 29551 --steps 20 --warmup 5 --sustain-s 2
OUT=r2_cfg2_n8.json run 8 29552 --workload cfg2 --steps 16 --warmup 3 --sustain-s 0
OUT=r2_scale_n4.json run 4 29553 --steps 20 --warmup 5 --sustain-s 0
OUT=r2_cfg2_n4.json run 4 29554 --workload cfg2 --steps 16 --warmup 3 --sustain-s 0
OUT=r2_scale_n2.json run 2 29555 --steps 20 --warmup 5 --sustain-s 0
OUT=r2_cfg2_n2.json run 2 29556 --workload cfg2 --steps 16 --warmup 3 --sustain-s 0
OUT=r2_scale_n1.json run 1 29557 --steps 20 --warmup 5 --sustain-s 0 --no-cpu-baseline
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T 200 $TR --master-port 29558 tools/config5.py --mode tiles --frames 8 --out gpurun_out/r2_config5_8k_n8_tiles.json > gpurun_out/cfg5a.log 2>&1; tail -1 gpurun_out/cfg5a.log | cut -c1-200
python - <<'PY'
import json
for f in ("r2_scale_n1", "r2_scale_n2", "r2_scale_n4", "r2_scale_n8", "r2_cfg2_n2", "r2_cfg2_n4", "r2_cfg2_n8"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d["bit_identity"], d["ms_gather_once"])
    except Exception as e:
        print(f, "unreadable", e)
PY
