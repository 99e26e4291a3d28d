#!/bin/bash
# First GPU call of round 2: validates what was built after round 1's GPU
# minutes were spent and measures the kNN row orders.  Everything lands in gpurun_out/.
#
#   gpurun --timeout 1500 -- 'bash tools/round2_first_call.sh'
#
# 1. the full GPU suite (must stay green: the defaults did not change)
# 2. the gated tests of the ordered kNN layout
# 3. tools/knn_bench.py at config 5 for each row order (slot and blocked layouts), with and without
#    evict-first slot loads
# 4. one ncu --set full capture of knn_fused_kernel under the Morton order (only after 3 exited 0)
set -u
out=gpurun_out
mkdir -p "$out"
python -m pytest tests -m gpu -x -q > "$out/r2_gpu_tests.log" 2>&1
echo "gpu suite exit $?" | tee "$out/r2_status.txt"
GL_RUN_UNVALIDATED=1 python -m pytest tests/test_knn_ordering.py -m gpu -q > "$out/r2_knn_ordering_tests.log" 2>&1
echo "ordered-layout tests exit $?" | tee -a "$out/r2_status.txt"
: > "$out/r2_knn_order.jsonl"
ok=0
for args in "0 0" "0 1" "256 1" "1024 1" "4096 1" "-1 1" "256 1 blocked" "-1 1 blocked" "0 1 blocked"; do
  # shellcheck disable=SC2086
  timeout 300 python tools/knn_bench.py 1000000 20 200 $args >> "$out/r2_knn_order.jsonl" 2>> "$out/r2_knn_order.err" || ok=1
done
for args in "0 1" "256 1"; do
  # shellcheck disable=SC2086
  GL_KNN_STREAM=1 timeout 300 python tools/knn_bench.py 1000000 20 200 $args >> "$out/r2_knn_order_stream.jsonl" 2>> "$out/r2_knn_order.err" || ok=1
done
echo "knn_bench sweep exit $ok" | tee -a "$out/r2_status.txt"
if [ "$ok" = 0 ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:knn_fused_kernel -s 30 -c 1 \
    -o "$out/r2_knn_fused_morton" -f python tools/knn_bench.py 1000000 20 40 0 1 > "$out/r2_ncu_knn.log" 2>&1
  echo "ncu capture exit $?" | tee -a "$out/r2_status.txt"
fi
tail -3 "$out/r2_gpu_tests.log" "$out/r2_knn_ordering_tests.log"
cat "$out/r2_knn_order.jsonl" "$out/r2_knn_order_stream.jsonl" 2>/dev/null | cut -c1-400
