#!/usr/bin/env bash
# N-GPU scaling checks (N = first argument, default 8): exchange probe, then bench with the NCCL and the NVLS exchange
N=${1:-8}
NARG=$N
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
[ -n "$SKIP_PROBE" ] || $T --master-port 29631 scripts/allreduce_probe2.py 2>&1 | grep -v "^\*\|OMP_NUM\|^$" | tail -12 | tee gpurun_out/allreduce_probe_n$N.txt
for cfg in "nccl event" "nvls event" "nvls split"; do
  set -- $cfg; ex=$1; md=$2
  PB_B200_DDP_MODE=$md PB_B200_EXCHANGE=$ex $T --master-port 29641 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_${ex}_$md.json 2> gpurun_out/bench_n${N}_${ex}_$md.err
  python -c "
import json;d=json.loads([l for l in open('gpurun_out/bench_n${N}_${ex}_$md.json') if l.startswith('{')][-1]);print('N=$N', '$ex', '$md', round(d['value']), round(d['ms_per_step'],4), 'ddp_check', d.get('ddp_check'), {k:d.get(k) for k in ('exposed_comm_us','compute_only_us','allreduce_alone_us')})" || tail -5 gpurun_out/bench_n${N}_${ex}_$md.err
done
python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-gpu-baseline --no-kernel-table > gpurun_out/bench_n1.json 2>/dev/null
python -c "
import json;d=json.load(open('gpurun_out/bench_n1.json'));print('N=1', round(d['value']), round(d['ms_per_step'],4))"
