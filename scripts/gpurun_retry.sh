#!/usr/bin/env bash
# gpurun_retry.sh TIMEOUT 'command' [extra gpurun flags...]: retry while the pod answers "busy" (exit code 3)
T="$1"; CMD="$2"; shift 2
for i in $(seq 1 40); do
    /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD"
    rc=$?
    if [ "$rc" != "3" ]; then exit $rc; fi
    sleep 90
done
exit 3
