#!/bin/bash
# ncu capture of ONE thin-round evaluation launch (B = 1: one parallel-in-time request per CTA) with the source page,
# to see where the primal chain's cycles go.  Usage (GPU box, repo root): bash tools/gpu_profile_pit.sh
out=gpurun_out
tag=${TAG:-r02_pit}
export MPCB_LOOP=host PROF_B=1 PROF_REPS=1
mkdir -p $out
python tools/prof_solve.py > $out/${tag}_plain.log 2>&1 || { echo "plain run failed"; cat $out/${tag}_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_eval_requests -s ${SKIP:-61} -c 1 -f -o $out/${tag} \
  python tools/prof_solve.py > $out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $out/${tag}.ncu-rep --page raw --csv > $out/${tag}_raw.csv 2>/dev/null
ncu -i $out/${tag}.ncu-rep --page source --csv 2>/dev/null | gzip -9 > $out/${tag}_source.csv.gz
rm -f $out/${tag}.ncu-rep
ls -la $out | grep ${tag}
