#!/usr/bin/env bash
# round-end N-GPU bench lines (N = first argument): config 2 and config 5
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
for c in 2 5; do
  timeout 400 $T --master-port 2964$c bench.py --gpus $N --config $c --steps 30 --warmup 5 > gpurun_out/r02_bench_config${c}_n$N.json 2> gpurun_out/r02_bench_config${c}_n$N.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r02_bench_config${c}_n$N.json") if l.startswith("{")][-1])
    print("config $c N=$N", round(d["value"]), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), "ddp_check", d.get("ddp_check"), {k:d.get(k) for k in ("exposed_comm_us","compute_only_us","allreduce_alone_us","exchange","ddp_mode","exchanges_per_step")})
except Exception as e:
    print("config $c N=$N failed", e); print(open("gpurun_out/r02_bench_config${c}_n$N.err").read()[-1500:])
PY
done
