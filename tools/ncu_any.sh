#!/bin/bash
# Full ncu capture of one launch of a kernel in an arbitrary command, after a plain run of it.
#   gpurun -- 'bash tools/ncu_any.sh <kernel-regex> <skip> <out-name> <cmd...>'
set -u
K=$1; SKIP=$2; OUT=$3; shift 3
mkdir -p gpurun_out
"$@" > gpurun_out/${OUT}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f -o gpurun_out/$OUT "$@" > gpurun_out/${OUT}_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/${OUT}_ncu.log
