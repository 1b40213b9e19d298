#!/usr/bin/env bash
# N-GPU bench of the data-parallel step under several exchange settings: "EXCHANGE MODE CTAS THREADS OVERLAP" per line
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
while read -r ex md ctas thr ov; do
  [ -z "$ex" ] && continue
  tag="${ex}_${md}_${ctas}x${thr}_ov${ov}"
  PB_B200_DDP_MODE=$md PB_B200_EXCHANGE=$ex PB_B200_NVLS_CTAS=$ctas PB_B200_NVLS_THREADS=$thr $T --master-port 29641 bench.py --gpus $N --steps 30 --warmup 5 --ddp-overlap $ov > gpurun_out/bench_n${N}_$tag.json 2> gpurun_out/bench_n${N}_$tag.err
  python -c "
import json;d=json.loads([l for l in open('gpurun_out/bench_n${N}_$tag.json') if l.startswith('{')][-1]);print('N=$N', '$tag', round(d['value']), round(d['ms_per_step'],4), 'ddp_check', d.get('ddp_check'), {k:round(d.get(k),1) for k in ('exposed_comm_us','compute_only_us','allreduce_alone_us')})" || tail -5 gpurun_out/bench_n${N}_$tag.err
done
