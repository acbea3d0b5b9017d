#!/bin/bash
# One-off tuning probe (GPU box, repo root): window-policy thresholds and a 64-thread transition CTA at B = 65536.
export PROBE_REPS=2
for t in "4096,12288" "2048,8192" "8192,24576" "4096,32768" "1024,4096"; do
  MPCB_SPEC_LIVE=$t PROBE="rounds:8:65536" PROBE_OUT=x.json timeout 300 python tools/mode_probe.py 2>&1 | grep "^{" | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); p=r['profile']; print('live $t', round(r['ms'],1), 'rounds', p['rounds'], 'E', round(p['eval_ms']), 'T', round(p['trans_ms']), 'evals/solve', round(r['evals_value']+r['evals_grad']))"
done
[ -f build/libmpcb200_blk64.so ] && MPCB_LIB=$PWD/build/libmpcb200_blk64.so PROBE="rounds:8:65536,rounds:8:4096" PROBE_OUT=y.json timeout 300 python tools/mode_probe.py 2>&1 | grep "^{" | cut -c1-420
