#!/bin/bash
# Runs on the GPU box (under gpurun): Nsight Compute captures of tools/prof_run.py workloads, exported to CSV there so that
# only small files travel back (gpurun_out/ is limited to 64 MiB; the .ncu-rep files are deleted).
#   tools/ncu_box.sh <tag> <mode> [kernel regex] [extra prof_run args]
set -x
tag=$1; mode=$2; regex=${3:-k_}; shift 3
out=gpurun_out
python tools/prof_run.py $mode "$@" > $out/${tag}_plain_$mode.log 2>&1 || { echo "plain run failed"; tail -5 $out/${tag}_plain_$mode.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file $out/${tag}_launches_$mode.csv \
    python tools/prof_run.py $mode "$@" > $out/${tag}_ncu1_$mode.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:$regex -f -o /tmp/${tag}_full_$mode \
    python tools/prof_run.py $mode "$@" > $out/${tag}_ncu2_$mode.log 2>&1
echo "full $mode rc=$?"
ncu -i /tmp/${tag}_full_$mode.ncu-rep --page raw --csv > $out/${tag}_raw_$mode.csv 2>/dev/null
ncu -i /tmp/${tag}_full_$mode.ncu-rep --page source --csv --print-source sass > /tmp/${tag}_src_$mode.csv 2>/dev/null
# the SASS page of every captured launch is large: keep it gzipped
gzip -9 -c /tmp/${tag}_src_$mode.csv > $out/${tag}_src_$mode.csv.gz
ls -la $out /tmp/${tag}_full_$mode.ncu-rep
