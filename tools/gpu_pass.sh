#!/bin/bash
# One GPU pass: parity tests, smoke, bench (+table), ncu launch list and one --set full capture.
# Usage (from the repo root, on the GPU box): bash tools/gpu_pass.sh <tag>
TAG=${1:-pass}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
tail -3 $OUT/${TAG}_pytest.log
python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/${TAG}_smoke.log
python bench.py --table > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench_table.txt; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>$OUT/${TAG}_bench_ref.err; echo "ref rc=$?"
PROF="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --serial --profile-range"
$PROF > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 200 --csv --log-file $OUT/${TAG}_launches.csv $PROF > $OUT/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
$PROF > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    --profile-from-start off \
    -k 'regex:crop_kernel|group_point_vec4|three_interpolate_nlc_vec4|fps_cluster|fps_multi|nms_filter|nms_clip|nms_sweep|three_nn|query_ball|ball_query|cloud_index' \
    -c 40 -o $OUT/${TAG}_full $PROF > $OUT/${TAG}_ncu2.log 2>&1
echo "full capture rc=$?"
ls -la $OUT
