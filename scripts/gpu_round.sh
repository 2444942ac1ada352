#!/bin/bash
# One GPU-box pass for a round: the gpu tests, the bench line of both arms, ncu launch lists and full captures.
# usage (from the repo root, under gpurun):  bash scripts/gpu_round.sh <tag> [tests|bench|ncu ...]
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
tag=${1:-r2}; shift
what=${*:-tests bench ncu}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $out/${tag}_gpu.txt 2>&1
nproc >> $out/${tag}_gpu.txt
for w in $what; do
case $w in
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest.log
  tail -3 $out/${tag}_pytest.log ;;
bench)
  timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
  timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench.err; echo "ref rc=$?" ;;
ncu)
  M="--metrics gpu__time_duration.sum --clock-control none --csv"
  timeout 600 ncu $M -c 400 --log-file $out/${tag}_launches_c5.csv python scripts/run_c5_once.py c5 2 > $out/${tag}_ncu.log 2>&1
  timeout 600 ncu $M -c 400 --log-file $out/${tag}_launches_c2.csv python scripts/run_c2_once.py c2 2 >> $out/${tag}_ncu.log 2>&1
  F="--set full --clock-control none --import-source on"
  # C5: skip the first iteration's kernels, capture the second one
  n=$(grep -c '^"[0-9]' $out/${tag}_launches_c5.csv); per=$(( n / 2 ))
  timeout 900 ncu $F -s $per -c $per -f -o $out/${tag}_c5 python scripts/run_c5_once.py c5 2 >> $out/${tag}_ncu.log 2>&1
  timeout 900 ncu $F -k regex:"k_linearize2|k_schur_smem|k_solve6|k_backsub_fused|k_tables|k_volume_tc|k_corr_lookup|k_to_pixel" -c 14 -f -o $out/${tag}_c2 python scripts/run_c2_once.py c2 1 >> $out/${tag}_ncu.log 2>&1
  tail -3 $out/${tag}_ncu.log ;;
esac
done
