#!/bin/bash
# same-box A/B of libsbo.so variants kept under scripts/probe/ (timing only): ab.sh "orig g2 g3" "1024 4096"
L=safe_bayesian_optimization_b200/csrc/libsbo.so
cp $L /tmp/libsbo_keep.so
for rep in 1 2; do
for v in $1; do
  cp scripts/probe/libsbo_$v.so $L
  for n in $2; do echo -n "$v: "; python scripts/dbg_probe.py $n 0 | tail -1; done
done
done
cp /tmp/libsbo_keep.so $L
