#!/bin/bash
# Same-box A/B of two builds of the library: ab/libawd_<name>.so, alternated; prints sb_probe.py's lines.
# usage: profiles/ab.sh "<names>" "<streams list>" [rounds]
L=pytorch-zappa-serverless_b200/awd_b200/libawd_b200.so
cp $L /tmp/keep.so
for r in $(seq 1 ${3:-2}); do
  for n in $1; do
    cp ab/libawd_$n.so $L
    echo "== $n (round $r)"
    PROBE_BS=$2 PROBE_T=${PROBE_T:-128} timeout 150 python profiles/sb_probe.py 2>&1 | grep "^bs" | cut -c1-260
  done
done
cp /tmp/keep.so $L
