#!/bin/sh
# Builds the C restatement of the oracle (test infrastructure) -> oracle/_build/liboracle.so
set -e
here="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$here/_build"
gcc -O3 -march=x86-64-v2 -std=c11 -shared -fPIC -pthread -o "$here/_build/liboracle.so" "$here/segpath_c.c"
