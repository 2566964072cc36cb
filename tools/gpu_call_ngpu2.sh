#!/bin/bash
# r2 multi-GPU call after the e2e / pool changes: usage gpu_call_ngpu2.sh N
cd "$(dirname "$0")/.."
N=${1:-8}
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29561 bench.py --gpus $N --steps 20 --warmup 5 --cpu-seconds 0 > $O/r2_bench_n$N.json 2> $O/r2_bench_n$N.err; echo "bench rc=$?"
{ echo "# rtb_render_multi: ONE host call, host buffers, final_scene 3840x2160 (tools/multi_bench.py), $N x B200"; echo "## 512 spp"; python tools/multi_bench.py 3840 2160 512; echo "## 64 spp"; python tools/multi_bench.py 3840 2160 64 0 3; } > $O/r2_multi_call.txt 2>&1; cat $O/r2_multi_call.txt
python -m pytest tests/test_gpu_peer.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2
$TR --master-port 29562 tools/shard_modes_bench.py 64 10000 2>/dev/null > $O/r2_shard_modes_n$N.md; cat $O/r2_shard_modes_n$N.md
python - <<PY
import json
d = json.load(open("$O/r2_bench_n$N.json"))
print("value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1), d.get("exchange"))
PY
tail -4 $O/r2_bench_n$N.err
