#!/bin/bash
# build, then (only if the build succeeded) run the given command on the GPU box
set -e
make -C motionplanning_b200/csrc > /tmp/make.log 2>&1 || { grep -i "error" -A3 motionplanning_b200/csrc/build.log | head -30; exit 1; }
/usr/local/graft/bin/gpurun --timeout ${GPU_TIMEOUT:-900} -- "$@" 2>&1 | grep -v "^\[gpurun\] sending" | tail -${TAIL:-12}
