# compute-sanitizer passes over the small-shape GPU tests (memcheck: out-of-bounds / misaligned
# accesses; racecheck: shared-memory hazards; synccheck: barrier misuse).  Slow: run on a subset.
set -x
mkdir -p gpurun_out
SUB='edge_sizes or acq_reduce_bit_exact or predict_acq_fused or empty_inputs or potrf_trtri_mll or graph_capturable'
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "$SUB" > gpurun_out/sanitize_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/sanitize_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "edge_sizes or acq_reduce_bit_exact or predict_acq_fused" > gpurun_out/sanitize_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/sanitize_racecheck.log
timeout 600 compute-sanitizer --tool synccheck --error-exitcode 9 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "edge_sizes or predict_acq_fused" > gpurun_out/sanitize_synccheck.log 2>&1; echo "synccheck rc=$?" >> gpurun_out/sanitize_synccheck.log
tail -4 gpurun_out/sanitize_memcheck.log gpurun_out/sanitize_racecheck.log gpurun_out/sanitize_synccheck.log
