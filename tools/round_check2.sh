# Second measurement pass (1 GPU): stage timers for C3/C4/C5, ncu full capture of the n<=64 predictive kernel
set -x
mkdir -p gpurun_out
python tools/c3_sweep.py 125000 2000 64 0 > gpurun_out/c3_sweep.log 2>&1
python tools/c4_step.py > gpurun_out/c4_step.log 2>&1
python tools/c5_step.py > gpurun_out/c5_step.log 2>&1
python tools/c3_mgpu.py 1000000 2000 2 > gpurun_out/c3_1gpu.json 2> gpurun_out/c3_1gpu.err
timeout 600 ncu --set full --import-source on --clock-control none -f -k regex:predict_tiny -c 1 -o gpurun_out/ncu_predict_tiny python tools/c3_sweep.py 32768 2000 64 0 > gpurun_out/ncu5.log 2>&1
cat gpurun_out/c3_sweep.log gpurun_out/c4_step.log gpurun_out/c5_step.log gpurun_out/c3_1gpu.json; tail -3 gpurun_out/c3_1gpu.err
