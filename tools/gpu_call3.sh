#!/bin/bash
# r2 GPU call 3: BVH4 vs BVH2 A/B (+ two-steps-per-vote variant), parity of the new tree, sanitizer passes, ncu at the bench shape
cd "$(dirname "$0")/.."
O=gpurun_out
SC=final_scene,final4k,cornell_box,random_scene,cornell_smoke,two_perlin_sphere,simple_light
{
echo "== BVH4 (default)";            python tools/ab.py --stats --scenes $SC raytracing_rust_b200/librtb200.so
echo "== BVH2 (reserved[6]=2)";      python tools/ab.py --stats --scenes $SC --knobs 0,0,0,0,0,0,2 raytracing_rust_b200/librtb200.so
echo "== BVH4, two steps per vote";  python tools/ab.py --scenes $SC exp/unroll2/librtb200.so
echo "== BVH2, two steps per vote";  python tools/ab.py --scenes $SC --knobs 0,0,0,0,0,0,2 exp/unroll2/librtb200.so
echo "== BVH4 again (drift check)";  python tools/ab.py --scenes final_scene,final4k raytracing_rust_b200/librtb200.so
} > $O/r2_ab_bvh.txt 2>&1
python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh_build.py -m gpu -q -rP -k "four_wide or closest_hit or pool or dual or wavefront or kat or bvh" > $O/r2_gputest3.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_gputest3.log
for tool in memcheck racecheck initcheck synccheck; do
  echo "== compute-sanitizer --tool $tool" >> $O/r2_sanitizer.txt
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_target.py 2>&1 | grep -v "^$" | tail -25 >> $O/r2_sanitizer.txt
done
tail -5 $O/r2_sanitizer.txt
for v in bvh4 bvh2; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_mega -s 1 -c 1 -f -o $O/r2_mega_4k64_$v python tools/profile_target.py final_scene 3840 2160 64 1 $v > $O/r2_ncu_$v.log 2>&1
  python tools/ncu_summary.py $O/r2_mega_4k64_$v.ncu-rep > $O/r2_mega_4k64_${v}_summary.txt 2>&1
done
ls -la $O | tail -12
