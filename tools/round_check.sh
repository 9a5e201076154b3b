# Round check on one B200: GPU tests, smoke, bench, stage timers, ncu launch list + full captures.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python tools/chol_only.py 1000 2048 > gpurun_out/chol_only.log 2>&1
python tools/acq_only.py 2000 524288 > gpurun_out/acq_only.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
NCU="ncu --set full --import-source on --clock-control none -f"
timeout 600 $NCU -k regex:tile_kernel --launch-skip 15 --launch-count 2 -o gpurun_out/ncu_tile_potrf python tools/chol_only.py 592 2048 1 > gpurun_out/ncu1.log 2>&1
timeout 600 $NCU -k regex:tile_kernel --launch-skip 52 --launch-count 2 -o gpurun_out/ncu_tile_trtri python tools/chol_only.py 592 2048 1 > gpurun_out/ncu2.log 2>&1
timeout 600 $NCU -k "regex:gram_kernel|diag_kernel" -c 3 -o gpurun_out/ncu_gram_diag python tools/chol_only.py 592 2048 1 > gpurun_out/ncu3.log 2>&1
timeout 600 $NCU -k regex:acq_reduce -c 1 -o gpurun_out/ncu_acq python tools/acq_only.py 2000 262144 > gpurun_out/ncu4.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log | tail -2; cat gpurun_out/bench_n1.json gpurun_out/bench_ref.json gpurun_out/chol_only.log gpurun_out/acq_only.log
