#!/bin/bash
# BASELINE configurations at their 8-GPU size (GPU box with 8 GPUs, repo root): metric batch, config 3, config 4, config 5.
out=gpurun_out; n=${1:-8}
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29531 bench.py --gpus $n --steps 3 --warmup 3 --no-api-leg > $out/r02_scale${n}_bench.json 2> $out/r02_scale${n}_bench.err; echo "bench rc=$?"
run 29532 bench.py --gpus $n --config 3 --steps 2 --warmup 3 > $out/r02_scale${n}_c3.json 2> $out/r02_scale${n}_c3.err; echo "c3 rc=$?"
run 29533 bench.py --gpus $n --config 4 --steps 2 --warmup 3 > $out/r02_scale${n}_c4.json 2> $out/r02_scale${n}_c4.err; echo "c4 rc=$?"
run 29534 bench.py --gpus $n --config 5 --steps 3 --warmup 3 > $out/r02_scale${n}_c5.json 2> $out/r02_scale${n}_c5.err; echo "c5 rc=$?"
run 29535 tools/config4_shard_check.py > $out/r02_scale${n}_c4_shard.log 2>&1; echo "c4 shard rc=$?"
for f in bench c3 c4 c5; do tail -c 300 $out/r02_scale${n}_$f.err; done
