#!/bin/bash
# One GPU pass: parity tests, smoke, bench (+table), reference arm, ncu launch list and one --set full capture.
# Usage (from the repo root, on the GPU box): bash tools/gpu_pass.sh <tag> [stages]
#   stages: any of  test smoke bench ref ncu   (default: all)
TAG=${1:-pass}
STAGES=${2:-"test smoke bench ref ncu"}
OUT=gpurun_out
mkdir -p $OUT
python -c "import heterofusionrcnn_b200 as hf; print(hf.build_id())" > $OUT/${TAG}_build.txt 2>/dev/null
echo "build $(cat $OUT/${TAG}_build.txt)"
has() { [[ " $STAGES " == *" $1 "* ]]; }
if has test; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
  tail -3 $OUT/${TAG}_pytest.log
fi
if has smoke; then
  timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/${TAG}_smoke.log
fi
if has bench; then
  timeout 900 python bench.py --table > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench_table.txt; echo "bench rc=$?"
  tail -c 600 $OUT/${TAG}_bench_table.txt | tail -5
fi
if has ref; then
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>$OUT/${TAG}_bench_ref.err; echo "ref rc=$?"
fi
if has ncu; then
  PROF="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-config5 --no-side-legs --serial --profile-range"
  timeout 300 $PROF > $OUT/${TAG}_plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 200 --csv --log-file $OUT/${TAG}_launches.csv $PROF > $OUT/${TAG}_ncu1.log 2>&1
  echo "launch list rc=$?"
  timeout 300 $PROF > $OUT/${TAG}_plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on \
      --profile-from-start off \
      -k 'regex:crop_kernel|group_point|three_interpolate|fps_|nms_filter|nms_clip|nms_sweep|three_nn|query_ball|ball_query|cloud_index|sa_tail|fp_chain' \
      -c 48 -o $OUT/${TAG}_full $PROF > $OUT/${TAG}_ncu2.log 2>&1
  echo "full capture rc=$?"
fi
ls -la $OUT | tail -20
