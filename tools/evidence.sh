#!/bin/bash
# One GPU call that refreshes the evidence of a round: tests, bench (both arms), launch list, one ncu --set full capture
# per kernel family.  Usage (from the repo root, through gpurun): bash tools/evidence.sh <tag>
tag=${1:-r5}
out=gpurun_out
python -m pytest tests -m gpu -q > $out/${tag}_gpu_tests.log 2>&1; tail -2 $out/${tag}_gpu_tests.log
for what in c2 c5 tamols; do
  python tools/profile_target.py $what 3 > /dev/null 2>&1 || echo "profile target $what failed"
  ncu --set full --clock-control none --import-source on -k regex:"k_rollout|k_tamols_" -s 2 -c 1 -o $out/${tag}_$what -f python tools/profile_target.py $what 3 > $out/${tag}_ncu_$what.log 2>&1
  ncu -i $out/${tag}_$what.ncu-rep --page details > $out/${tag}_ncu_details_$what.txt
  ncu -i $out/${tag}_$what.ncu-rep --page raw --csv > $out/${tag}_ncu_raw_$what.csv
  ncu -i $out/${tag}_$what.ncu-rep --page source --csv --print-source=sass > $out/${tag}_source_$what.csv
  python tools/ncu_source_summary.py $out/${tag}_source_$what.csv ${tag}_$what > $out/${tag}_ncu_source_$what.txt 2>&1
  rm -f $out/${tag}_$what.ncu-rep $out/${tag}_source_$what.csv  # tens of MB each: gpurun_out travels back only below 64 MiB
done
cp $out/${tag}_ncu_raw_*.csv profiles/  # bench.py reads roofline.traffic from profiles/<tag>_ncu_raw_*.csv
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; python tools/bench_extract.py $out/${tag}_bench.json
python bench.py --impl reference --steps 200 --warmup 20 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err
python tools/fp32_envelope.py --gpu > $out/${tag}_fp32_envelope.txt 2>&1
tools/ubench > $out/${tag}_ubench.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_bench.csv python bench.py --steps 5 --warmup 3 --no-cpu --no-parity > $out/${tag}_ncu_launch.log 2>&1
ls -la $out | tail -30
