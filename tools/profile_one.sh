#!/bin/bash
# One ncu --set full capture of one kernel family, summarised: bash tools/profile_one.sh <tag> <c2|c5|tamols>
tag=$1; what=$2; out=gpurun_out
python tools/profile_target.py $what 3 > /dev/null 2>&1 || echo "profile target $what failed"
ncu --set full --clock-control none --import-source on -k regex:"k_rollout|k_tamols_" -s 2 -c 1 -o $out/${tag}_$what -f python tools/profile_target.py $what 3 > $out/${tag}_ncu_$what.log 2>&1
ncu -i $out/${tag}_$what.ncu-rep --page details > $out/${tag}_ncu_details_$what.txt
ncu -i $out/${tag}_$what.ncu-rep --page raw --csv > $out/${tag}_ncu_raw_$what.csv
ncu -i $out/${tag}_$what.ncu-rep --page source --csv --print-source=sass > $out/${tag}_source_$what.csv
python tools/ncu_source_summary.py $out/${tag}_source_$what.csv ${tag}_$what > $out/${tag}_ncu_source_$what.txt 2>&1
rm -f $out/${tag}_$what.ncu-rep $out/${tag}_source_$what.csv
