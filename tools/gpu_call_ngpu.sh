#!/bin/bash
# r2 multi-GPU call: usage gpu_call_ngpu.sh N   (run under `gpurun --gpus N`)
cd "$(dirname "$0")/.."
N=${1:-8}
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 --cpu-seconds 0 > $O/r2_bench_n$N.json 2> $O/r2_bench_n$N.err; echo "bench default rc=$?"
$TR --master-port 29522 tools/shard_modes_bench.py 64 10000 > $O/r2_shard_modes_n$N.md 2> $O/r2_shard_modes_n$N.err; echo "shard modes rc=$?"; cat $O/r2_shard_modes_n$N.md
$TR --master-port 29523 bench.py --gpus $N --steps 10 --warmup 3 --cpu-seconds 0 --reduce nccl > $O/r2_bench_n${N}_nccl.json 2> $O/r2_bench_n${N}_nccl.err; echo "bench nccl rc=$?"
python -m pytest tests/test_gpu_peer.py -m gpu -q > $O/r2_gputest_peer_n$N.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r2_gputest_peer_n$N.log
for f in $O/r2_bench_n$N.json $O/r2_bench_n${N}_nccl.json; do python - "$f" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[1], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1), "exchange", d.get("exchange"))
PY
done
tail -5 $O/r2_bench_n$N.err
