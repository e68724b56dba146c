#!/bin/bash
# One GPU box, one GPU: the evidence the round's docs cite, taken from the tree as it stands.
#   gpurun --timeout 2400 -- 'bash tools/round_evidence.sh r02k'
# Writes gpurun_out/<tag>_*: pytest log, bench lines (both arms), the ncu launch list of the bench command, one
# `--set full` capture of the extraction kernels and of the matching kernels, and their summaries.
tag=${1:-r02k}
out=gpurun_out
mkdir -p $out
export PYTHONUNBUFFERED=1

timeout 1500 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest.txt 2>&1
tail -3 $out/${tag}_pytest.txt
# written in round 2 after the GPU budget was spent: the shim's object layer on the reference's real objects (gated until seen green)
ORBFE_TEST_REAL_OBJECTS=1 timeout 300 python -m pytest tests/test_shim_real_objects.py -m gpu -q > $out/${tag}_pytest_real_objects.txt 2>&1
tail -3 $out/${tag}_pytest_real_objects.txt

timeout 600 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err
tail -c 600 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err
tail -c 300 $out/${tag}_bench_reference.json

# launch list of the same (extraction + matching) command, no CPU baseline, no extra legs
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-legs > $out/${tag}_ncu_launches.log 2>&1
python tools/summarize_launches.py $out/${tag}_launches.csv > $out/${tag}_launches_summary.txt 2>&1
cat $out/${tag}_launches_summary.txt

# --set full: extraction kernels of one step (skip the warm-up launches), then the matching kernels
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:'k_fast|k_pyr|k_describe|k_blur|k_octree|k_assemble' --launch-skip 42 --launch-count 14 \
    -f -o $out/${tag}_extract python bench.py --steps 1 --warmup 3 --no-cpu --no-match --no-legs > $out/${tag}_ncu_extract.log 2>&1
ncu -i $out/${tag}_extract.ncu-rep --page raw --csv > $out/${tag}_extract_raw.csv 2>/dev/null
python tools/summarize_ncu_full.py $out/${tag}_extract_raw.csv $out/${tag}_extract_summary.json > $out/${tag}_extract_summary.txt 2>&1
cat $out/${tag}_extract_summary.txt

[ "$2" = noknn2 ] && exit 0     # the matching kernels did not change since the last capture
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:'k_knn2' --launch-skip 6 --launch-count 4 \
    -f -o $out/${tag}_knn2 python bench.py --steps 1 --warmup 3 --no-cpu --no-legs > $out/${tag}_ncu_knn2.log 2>&1
ncu -i $out/${tag}_knn2.ncu-rep --page raw --csv > $out/${tag}_knn2_raw.csv 2>/dev/null
python tools/summarize_ncu_full.py $out/${tag}_knn2_raw.csv $out/${tag}_knn2_summary.json > $out/${tag}_knn2_summary.txt 2>&1
cat $out/${tag}_knn2_summary.txt
