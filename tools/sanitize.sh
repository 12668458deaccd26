#!/bin/bash
# compute-sanitizer passes over the C-ABI kernels (run on the GPU box):  tools/sanitize.sh <tag>
# memcheck with torch's caching allocator off (every tensor its own cudaMalloc, so an out-of-bounds access leaves
# its allocation), then racecheck and initcheck on smoke().  Logs under gpurun_out/sanitize_<tag>/.
tag=${1:-run}
out=gpurun_out/sanitize_$tag
mkdir -p "$out"
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
san=/usr/local/cuda/bin/compute-sanitizer
run() { # name, limit seconds, tool, command...
    name=$1; limit=$2; tool=$3; shift 3
    timeout "$limit" $san --tool "$tool" --error-exitcode 9 --print-limit 20 "$@" > "$out/$name.log" 2>&1
    rc=$?
    echo "$name tool=$tool rc=$rc $(grep -c 'Invalid\|Race\|Uninitialized\|hazard' "$out/$name.log") findings; $(grep 'ERROR SUMMARY\|RACECHECK SUMMARY' "$out/$name.log" | tail -1); $(grep -E 'passed|failed' "$out/$name.log" | tail -1)" | tee -a "$out/summary.txt"
}
run smoke_memcheck 240 memcheck python -c "import __graft_entry__ as g; g.smoke()"
run smoke_racecheck 240 racecheck python -c "import __graft_entry__ as g; g.smoke()"
run lidar_memcheck 150 memcheck python -m pytest tests/test_gpu_lidar_input.py -x -q
run glue_memcheck 240 memcheck python -m pytest tests/test_gpu_glue.py -x -q
run knn_memcheck 240 memcheck python -m pytest tests/test_gpu_knn.py -x -q
run backward_memcheck 300 memcheck python -m pytest tests/test_gpu_backward.py -x -q
run parity_memcheck 480 memcheck python -m pytest tests/test_gpu_parity.py -x -q -k "fps_matches_oracle or query_ball_point_matches_oracle or pruned_matches_oracle or three_nn_matches_oracle or group_point or three_interpolate or crop_matches_oracle or crop_with_cloud_index or bev_iou_known or oriented_nms_matches_oracle or three_nn_weights or config0"
cat "$out/summary.txt"
