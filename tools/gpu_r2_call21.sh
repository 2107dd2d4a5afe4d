# round 2, call 21 (2 GPUs): N = 2 after the pipeline-count rule (sanity of the driver's N > 1 arm)
T="timeout -k 10"
run() { N=$1; PORT=$2; shift 2
  $T 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/tmp.log 2> gpurun_out/tmp.err
  tail -2 gpurun_out/tmp.err; grep '^{' gpurun_out/tmp.log > gpurun_out/$OUT; cut -c1-160 gpurun_out/$OUT; }
OUT=r2_scale_n2_final.json run 2 29561 --steps 20 --warmup 5 --sustain-s 0
OUT=r2_cfg2_n2_final.json run 2 29562 --workload cfg2 --steps 16 --warmup 3 --sustain-s 0
python - <<'PY'
import json
for f in ("r2_scale_n2_final", "r2_cfg2_n2_final"):
    d = json.load(open(f"gpurun_out/{f}.json"))
    print(f, round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d["bit_identity"])
PY
