#!/usr/bin/env bash
# round-end single-GPU evidence: tests, smoke, bench lines of configs 2 / 4 / 5, reference arm, profiles
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python bench.py > gpurun_out/r02_bench_config2.json 2> gpurun_out/r02_bench_config2.err
python bench.py --config 4 --no-gpu-baseline --no-cpu-baseline > gpurun_out/r02_bench_config4.json 2> gpurun_out/r02_bench_config4.err
python bench.py --config 5 --no-gpu-baseline --no-cpu-baseline > gpurun_out/r02_bench_config5.json 2> gpurun_out/r02_bench_config5.err
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err
bash scripts/capture_profiles.sh
