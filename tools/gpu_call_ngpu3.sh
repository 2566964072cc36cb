#!/bin/bash
# r2 multi-GPU confirmation at HEAD (kernel signature changed with FlatTables): usage gpu_call_ngpu3.sh N
cd "$(dirname "$0")/.."
N=${1:-8}
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
python -m pytest tests/test_gpu_peer.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2
$TR --master-port 29571 bench.py --gpus $N --steps 10 --warmup 3 --cpu-seconds 0 > $O/r2_bench_n${N}_head.json 2> $O/r2_bench_n${N}_head.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("$O/r2_bench_n${N}_head.json"))
print("value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1), d.get("exchange"))
PY
{ echo "## cornell_box 1024x1024 x 512 spp, one rtb_render_multi call on $N GPUs (flat scan on every shard)"; python - <<PY
import sys, time
sys.path.insert(0, ".")
import numpy as np, raytracing_rust_b200 as rtb
s, w, cam = rtb.Scene.named("cornell_box", 1024, 1024)
for rep in range(3):
    t = time.perf_counter()
    a, r8, st = s.render_multi(w, cam, 1024, 1024, 512, want_stats=True)
    dt = time.perf_counter() - t
    print(f"call {rep}: {dt*1e3:.1f} ms wall, {st['paths']/dt/1e6:.0f} Mpaths/s, mean {float(a.mean())/512:.4f}")
s1, w1, c1 = rtb.Scene.named("cornell_box", 1024, 1024)
b, _, _ = s1.render(w1, c1, 1024, 1024, 512)
print("max |multi - single| / 512 =", float(np.abs(a - b).max()) / 512)
PY
} 2>&1 | tee $O/r2_multi_flat_n$N.txt
