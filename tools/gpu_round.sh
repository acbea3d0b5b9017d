#!/bin/bash
# One consolidated GPU pass: parity tests, smoke, both bench arms, then the ncu launch list of the bench command.
# Usage (from the repo root on the GPU box): bash tools/gpu_round.sh <tag>
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${tag}_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "ref rc=$?"
python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; rc=$?; echo "bench rc=$rc"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_under_ncu.log 2>&1; echo "ncu rc=$?"
fi
tail -3 $out/${tag}_pytest_gpu.log; cat $out/${tag}_smoke.log | tail -2; cat $out/${tag}_bench_reference.json; cat $out/${tag}_bench.json
