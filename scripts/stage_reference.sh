#!/usr/bin/env bash
# Stage the reference's own hot-path files next to the benchmark so that the CPU baseline (bench.py --impl reference,
# the cpu_baseline leg) and the eager-torch-on-B200 rows time the UNMODIFIED reference modules, not a restatement.
#
#   scripts/stage_reference.sh [/path/to/reference]      (default /root/reference)
#
# The copies go to baseline/_ref/, which is git-ignored (reference sources never enter this repository's history)
# but not gpurun-ignored, so they travel to the GPU box with the working tree.  bench.py falls back to the oracle
# port (cpu_baseline.kind = "port") when the directory is absent.
set -euo pipefail
SRC="${1:-/root/reference}"
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
DST="$ROOT/baseline/_ref"
mkdir -p "$DST"
for f in modules.py utils.py data_processing.py dummy_tests.py; do
    cp "$SRC/ProteinBERT/$f" "$DST/$f"
done
( cd "$DST" && sha256sum modules.py utils.py data_processing.py dummy_tests.py > SHA256SUMS )
echo "staged $(ls "$DST" | wc -l) files in $DST"
