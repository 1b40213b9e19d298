#!/usr/bin/env bash
# 2-GPU checks: gradient parity across ranks, then bench at N=2 with and without the overlapped exchange
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
mkdir -p gpurun_out
$T --master-port 29611 scripts/ddp_check.py --out gpurun_out/ddp_check_fp32.json 2>&1 | tail -3
$T --master-port 29612 scripts/ddp_check.py --precision bf16 --out gpurun_out/ddp_check_bf16.json 2>&1 | tail -2
for ov in 1 0; do
  $T --master-port 2962$ov bench.py --gpus 2 --steps 30 --warmup 5 --ddp-overlap $ov > gpurun_out/bench_n2_ov$ov.json 2> gpurun_out/bench_n2_ov$ov.err
  python -c "
import json;d=json.loads([l for l in open('gpurun_out/bench_n2_ov$ov.json') if l.startswith('{')][-1]);print('N=2 overlap=$ov', round(d['value']), round(d['ms_per_step'],4), 'ddp_check', d.get('ddp_check'), {k:d.get(k) for k in ('exposed_comm_us','compute_only_us','allreduce_alone_us','bucket_loose_tensors')})" || tail -5 gpurun_out/bench_n2_ov$ov.err
done
python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-gpu-baseline --no-kernel-table > gpurun_out/bench_n1.json 2>/dev/null
python -c "
import json;d=json.load(open('gpurun_out/bench_n1.json'));print('N=1', round(d['value']), round(d['ms_per_step'],4))"
