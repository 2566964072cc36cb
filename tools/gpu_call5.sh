#!/bin/bash
# r2 GPU call 5 (end of round, HEAD): the whole -m gpu suite with the parity table, the N=1 bench line, the reference arm
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests -m gpu -q -rP > $O/r2_gputest5.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_gputest5.log
python tools/parity_report.py $O/r2_gputest5.log > $O/r2_parity5.md
python bench.py --steps 10 --warmup 3 > $O/r2_bench_n1_head.json 2> $O/r2_bench_n1_head.err; echo "bench rc=$?"; cut -c1-300 $O/r2_bench_n1_head.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_bench_ref_head.json 2> $O/r2_bench_ref_head.err; echo "ref rc=$?"; cut -c1-300 $O/r2_bench_ref_head.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
