#!/bin/bash
# Round-2 ncu evidence (run under gpurun on one B200; every profiled command first exits 0 without ncu).
#   launch list of the bench command, then `--set full` captures of the persistent generation kernel at 1024 and at 128 streams.
# Raw exports land in gpurun_out/; profiles/ncu_traffic.py turns them into the committed csv / json under profiles/.
set -x
B="python bench.py --steps 1 --warmup 3 --new-tokens 8 --no-cpu-baseline --no-latency"
$B > gpurun_out/plain_1024.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_1024.csv $B > gpurun_out/ncu_l.log 2>&1
$B > gpurun_out/plain_1024b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mega_generate -s 3 -c 1 -o gpurun_out/r02_mega_1024 $B > gpurun_out/ncu_f1.log 2>&1
$B --streams 128 > gpurun_out/plain_128.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mega_generate -s 3 -c 1 -o gpurun_out/r02_mega_128 $B --streams 128 > gpurun_out/ncu_f2.log 2>&1
for n in 1024 128; do
  ncu -i gpurun_out/r02_mega_$n.ncu-rep --page raw --csv > gpurun_out/r02_mega_${n}_raw.csv 2>/dev/null
  ncu -i gpurun_out/r02_mega_$n.ncu-rep --page details --csv > gpurun_out/r02_mega_${n}_details.csv 2>/dev/null
done
ls -la gpurun_out/
