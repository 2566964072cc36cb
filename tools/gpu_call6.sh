#!/bin/bash
# r2 GPU call 6 (HEAD, after the block-sized shared arrays of E17): the self-checking build over every kernel + the random-size soak
cd "$(dirname "$0")/.."
O=gpurun_out
{
sed -n 1,8p profiles/r2_sanitizer.txt
echo "== RTB200_LIB=exp/checks/librtb200.so python tools/sanitize_target.py"
RTB200_LIB=exp/checks/librtb200.so timeout 600 python tools/sanitize_target.py 2>&1 | tail -40
echo
echo "== RTB200_LIB=exp/checks/librtb200.so python tools/fuzz_sizes.py 90 11"
RTB200_LIB=exp/checks/librtb200.so timeout 300 python tools/fuzz_sizes.py 90 11 2>&1 | tail -5
} > $O/r2_sanitizer_head.txt 2>&1
tail -6 $O/r2_sanitizer_head.txt
