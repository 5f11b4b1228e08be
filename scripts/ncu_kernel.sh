#!/bin/bash
# usage: scripts/ncu_kernel.sh <tag> <kernel-regex> <launch-skip> -- <command...>
# One `ncu --set full` capture of one launch of one kernel -> gpurun_out/<tag>.ncu-rep
tag=$1; regex=$2; skip=$3; shift 4
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k "regex:$regex" --launch-skip "$skip" \
    --launch-count 1 -f -o "gpurun_out/$tag" "$@" > "gpurun_out/$tag.log" 2>&1
tail -3 "gpurun_out/$tag.log"
