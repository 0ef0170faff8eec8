# compute-sanitizer pass: memcheck over the hand-rolled mbarrier / TMA / TMEM kernels' tests,
# racecheck + synccheck over a bounded subset (racecheck serialises shared-memory traffic: ~50x slower).
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
export GM_PDL=0   # programmatic dependent launch is not modelled by the sanitizer's stream-order checks
for t in tests/test_gpu_geo.py tests/test_gpu_geo_tc.py tests/test_gpu_knn.py; do
  n=$(basename $t .py)
  timeout 900 $CS --tool memcheck --error-exitcode 7 --print-limit 20 python -m pytest $t -m gpu -x -q > gpurun_out/sanitizer_memcheck_$n.log 2>&1
  echo "memcheck $n rc=$?" | tee -a gpurun_out/sanitizer_summary.txt
done
timeout 900 $CS --tool racecheck --racecheck-report analysis --error-exitcode 7 --print-limit 20 python -m pytest tests/test_gpu_geo.py -m gpu -x -q -k "bit_identical or deterministic" > gpurun_out/sanitizer_racecheck_geo.log 2>&1
echo "racecheck geo rc=$?" | tee -a gpurun_out/sanitizer_summary.txt
timeout 900 $CS --tool racecheck --racecheck-report analysis --error-exitcode 7 --print-limit 20 python -m pytest tests/test_gpu_knn.py -m gpu -x -q -k "exact" > gpurun_out/sanitizer_racecheck_knn.log 2>&1
echo "racecheck knn rc=$?" | tee -a gpurun_out/sanitizer_summary.txt
timeout 600 $CS --tool synccheck --error-exitcode 7 --print-limit 20 python -m pytest tests/test_gpu_geo_tc.py -m gpu -x -q > gpurun_out/sanitizer_synccheck_geo_tc.log 2>&1
echo "synccheck geo_tc rc=$?" | tee -a gpurun_out/sanitizer_summary.txt
for f in gpurun_out/sanitizer_*.log; do echo "== $f"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" $f | tail -4; done
