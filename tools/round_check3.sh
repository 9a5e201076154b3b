# Headline refresh (1 GPU): bench, launch list, ncu full capture of the predictive kernel at the bench shape
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -f -k regex:predict_ws -c 1 -o gpurun_out/ncu_predict_ws_S1000 python tools/predict_only.py 1000 2048 100 > gpurun_out/ncu8.log 2>&1
cat gpurun_out/bench_n1.json; tail -1 gpurun_out/ncu8.log
