#!/usr/bin/env bash
# build the extension, then run a command on the GPU box:  tools/grun.sh <timeout_s> '<command>'
set -e
cd "$(dirname "$0")/.."
bash auditory-slow-fast_b200/build.sh 2>&1 | grep -E "error|built" || { echo "BUILD FAILED"; exit 1; }
to=$1; shift
/usr/local/graft/bin/gpurun --timeout "$to" -- "$@"
