#!/usr/bin/env bash
# build a probe library in a scratch copy (the shipped .so stays untouched) and run a probe script against it
set -e
DEFS="$1"; shift
TMP=$(mktemp -d)
cp -r proteinbert-pytorch-replication_b200 proteinbert_b200.py bench.py scripts include oracle "$TMP"/
cd "$TMP"
rm -rf proteinbert-pytorch-replication_b200/build proteinbert-pytorch-replication_b200/libpb_b200.so
PB_NVCC_DEFINES="$DEFS" python -c "import importlib; importlib.import_module('proteinbert-pytorch-replication_b200.build').build(force=True)"
"$@"
