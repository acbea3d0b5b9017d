#!/bin/bash
# Round-2 ncu pass (ONE gpurun call, ncu only): launch list of the bench command + `--set full` captures of the two
# round-mode kernels in the bulk and in the tail of a 64K batch.  Every profiled command first exits 0 without ncu.
# Usage (GPU box, repo root): bash tools/gpu_profile_r02.sh
out=gpurun_out
tag=r02
export MPCB_LOOP=host   # ncu profiles the launches of the host-driven loop (same kernels as the graph-driven default)
mkdir -p $out
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-api-leg"
$BENCH > $out/${tag}_bench_plain.json 2> $out/${tag}_bench_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/${tag}_launches.csv \
  $BENCH > $out/${tag}_bench_under_ncu.log 2>&1; echo "launch list rc=$?"
export PROF_B=65536 PROF_REPS=1
python tools/prof_solve.py > $out/${tag}_prof_plain.log 2>&1; rc=$?; echo "plain rc=$rc"; cat $out/${tag}_prof_plain.log
if [ $rc -eq 0 ]; then
  for spec in "k_eval_requests:100:eval_bulk" "k_solve:100:trans_bulk" "k_eval_requests:1500:eval_tail" "k_solve:1500:trans_tail"; do
    IFS=: read kern skip name <<< "$spec"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c 1 -f -o $out/${tag}_$name \
      python tools/prof_solve.py > $out/${tag}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
    # gpurun_out/ is capped at 64 MiB and a report with sources is ~20 MB: keep the CSV pages, drop the report
    ncu -i $out/${tag}_$name.ncu-rep --page raw --csv > $out/${tag}_${name}_raw.csv 2>/dev/null
    ncu -i $out/${tag}_$name.ncu-rep --page source --csv 2>/dev/null | gzip -9 > $out/${tag}_${name}_source.csv.gz
    rm -f $out/${tag}_$name.ncu-rep
  done
fi
ls -la $out | grep ${tag}_
