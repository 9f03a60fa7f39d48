#!/bin/bash
# ncu --set full of one scaled-scene render kernel (the timed launch of the second repetition), summarised on
# the box: tools/run_ncu.sh <tag> <n_objects> <w> <h> <depth> [mode] [bench workload name]     -> gpurun_out/<tag>.txt
# (+ gpurun_out/<tag>.capture.json when a workload name is given: what bench.py may quote, with the source hash)
set -u
TAG=$1; N=$2; W=$3; H=$4; D=$5; M=${6:-default}; WL=${7:-}
O=gpurun_out; mkdir -p $O
python tools/prof_scaled.py $N $W $H $D 2 $M > /dev/null || exit 1     # must exit 0 without ncu first
ncu --set full --clock-control none --import-source on -k regex:render_ -s 2 -c 1 -f -o /tmp/$TAG \
    python tools/prof_scaled.py $N $W $H $D 2 $M > $O/${TAG}_ncu.log 2>&1
{ python tools/ncu_summary.py /tmp/$TAG.ncu-rep; echo; echo "== executed code footprint (tools/ncu_footprint.py) =="; python tools/ncu_footprint.py /tmp/$TAG.ncu-rep;
  echo; echo "== hottest source lines (ncu --page source; share of samples / of executed instructions) =="; python tools/ncu_lines.py /tmp/$TAG.ncu-rep 45; } > $O/$TAG.txt 2>&1
python tools/ncu_lines.py /tmp/$TAG.ncu-rep 2000 > $O/${TAG}_all_lines.txt 2>&1
ncu -i /tmp/$TAG.ncu-rep --page source --csv --print-source sass > $O/${TAG}_sass.csv 2>/dev/null
[ -n "$WL" ] && python tools/ncu_capture_json.py /tmp/$TAG.ncu-rep $WL $TAG > /dev/null
head -8 $O/$TAG.txt
