#!/bin/bash
# ncu capture of ONE transition launch of a B = 1 solve (one tile) with the source page: where a lone tile's 38 us go.
# Usage (GPU box, repo root): bash tools/gpu_profile_trans1.sh
out=gpurun_out
tag=${TAG:-r02_trans1}
export MPCB_LOOP=host PROF_B=${PROF_B:-1} PROF_REPS=1
mkdir -p $out
python tools/prof_solve.py > $out/${tag}_plain.log 2>&1 || { echo "plain run failed"; cat $out/${tag}_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on --warp-sampling-interval 0 -k regex:k_solve -s ${SKIP:-61} -c 1 -f -o $out/${tag} \
  python tools/prof_solve.py > $out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $out/${tag}.ncu-rep --page raw --csv > $out/${tag}_raw.csv 2>/dev/null
ncu -i $out/${tag}.ncu-rep --page source --csv 2>/dev/null | gzip -9 > $out/${tag}_source.csv.gz
ncu -i $out/${tag}.ncu-rep --page source --print-source cuda --csv 2>/dev/null | gzip -9 > $out/${tag}_source_cuda.csv.gz
rm -f $out/${tag}.ncu-rep
ls -la $out | grep ${tag}
