#!/bin/bash
# gpu_round.sh + one `ncu --set full` capture of k_solve and of k_value_grad (each after its plain run exited 0).
# Usage (GPU box, repo root): bash tools/gpu_round_full.sh <tag>
tag=${1:-r01}
out=gpurun_out
bash tools/gpu_round.sh $tag
python tools/prof_solve.py > $out/${tag}_prof_solve_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve -s 1 -c 1 -f -o $out/${tag}_k_solve \
  python tools/prof_solve.py > $out/${tag}_ncu_solve.log 2>&1; echo "ncu solve rc=$?"
PROF_B=262144 python tools/prof_eval.py > $out/${tag}_prof_eval_plain.log 2>&1 && \
PROF_B=262144 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_value_grad -s 1 -c 1 -f -o $out/${tag}_k_value_grad \
  python tools/prof_eval.py > $out/${tag}_ncu_eval.log 2>&1; echo "ncu eval rc=$?"
cat $out/${tag}_prof_solve_plain.log
ls -la $out
