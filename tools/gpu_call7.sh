#!/bin/bash
# r2 GPU call 7: the flat scan (E19, E20) as the default on small scenes: whole -m gpu suite, self-check build, tree A/B with the automatic choice, configs
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests -m gpu -q -rP > $O/r2_gputest7.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_gputest7.log
python tools/parity_report.py $O/r2_gputest7.log > $O/r2_parity7.md
{
sed -n 1,8p profiles/r2_sanitizer.txt
echo "== RTB200_LIB=exp/checks/librtb200.so python tools/sanitize_target.py"
RTB200_LIB=exp/checks/librtb200.so timeout 600 python tools/sanitize_target.py 2>&1 | tail -44
echo
echo "== RTB200_LIB=exp/checks/librtb200.so python tools/fuzz_sizes.py 60 13"
RTB200_LIB=exp/checks/librtb200.so timeout 300 python tools/fuzz_sizes.py 60 13 2>&1 | tail -5
} > $O/r2_sanitizer7.txt 2>&1
tail -4 $O/r2_sanitizer7.txt
python tools/tree_ab.py > $O/e19_tree_ab2.txt 2>&1; cat $O/e19_tree_ab2.txt
python bench.py --configs-table --cpu-seconds 6 > $O/r2_configs7.md 2> $O/r2_configs7.err; echo "configs rc=$?"; cat $O/r2_configs7.md
