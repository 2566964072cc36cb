#!/bin/bash
# r2 GPU call 4: the self-checking build over every kernel (stand-in for compute-sanitizer), the whole -m gpu suite with the
# parity table, the N=1 bench line, the five-config table
cd "$(dirname "$0")/.."
O=gpurun_out
{
echo "# r2 sanitizer record"
echo "compute-sanitizer (memcheck / racecheck / initcheck / synccheck) is CLOSED on this GPU pool; gpurun answered, for each tool:"
echo "  'compute-sanitizer is closed on this pool and stays closed: runs under it have left GPUs needing a reset. Find a bad access"
echo "   with bounds checks and asserts of your own, small cases, and a comparison with the CPU reference'"
echo "Stand-in: the -DRTB_DEBUG_CHECKS=1 build (every stack / node / primitive / material / texture / texel / tile / pixel / queue index"
echo "bounds-checked on the device, NaN radiance counted), run over every kernel (tools/sanitize_target.py) and under the"
echo "random-size soak (tools/fuzz_sizes.py: sizes, sample + row-band shards, both trees, both pipelines, occupancy points)."
echo
echo "== RTB200_LIB=exp/checks/librtb200.so python tools/sanitize_target.py"
RTB200_LIB=exp/checks/librtb200.so timeout 600 python tools/sanitize_target.py 2>&1 | tail -40
echo
echo "== RTB200_LIB=exp/checks/librtb200.so python tools/fuzz_sizes.py 60 7"
RTB200_LIB=exp/checks/librtb200.so timeout 300 python tools/fuzz_sizes.py 60 7 2>&1 | tail -5
} > $O/r2_sanitizer.txt 2>&1
tail -4 $O/r2_sanitizer.txt
python -m pytest tests -m gpu -q -rP > $O/r2_gputest4.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_gputest4.log
python tools/parity_report.py $O/r2_gputest4.log > $O/r2_parity.md
python bench.py --steps 10 --warmup 3 > $O/r2_bench_n1.json 2> $O/r2_bench_n1.err; echo "bench rc=$?"; cut -c1-400 $O/r2_bench_n1.json
python bench.py --configs-table --cpu-seconds 6 > $O/r2_configs.md 2> $O/r2_configs.err; echo "configs rc=$?"; cat $O/r2_configs.md
