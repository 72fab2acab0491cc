#!/usr/bin/env bash
# TEST INFRASTRUCTURE. Builds the plain-C restatement oracle/rt_port.c -> oracle/_build/librt_port.so
# (git-ignored; travels to the GPU box with the snapshot, and can be rebuilt there: gcc is in the image).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$HERE/_build"
if [ "$HERE/_build/librt_port.so" -nt "$HERE/rt_port.c" ]; then exit 0; fi
gcc -std=c11 -O2 -ffp-contract=off -fno-fast-math -fPIC -fopenmp -fvisibility=hidden -Wall -Wno-unused-function \
    -shared "$HERE/rt_port.c" -o "$HERE/_build/librt_port.so" -lm
echo "built $HERE/_build/librt_port.so"
