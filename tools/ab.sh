#!/bin/sh
# A/B two builds of libssb200.so in the same GPU session (tuning aid).
for rep in 1 2; do
  for lib in "$@"; do
    echo "== $lib (rep $rep)"
    SSB200_LIB=$lib SWEEP_CFGS="${AB_CFGS:-4,4 8,3}" SWEEP_WAVES=1 python tools/sweep_argmax.py | grep cfg
  done
done
