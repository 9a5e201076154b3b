set -x
python tools/chol_only.py 1000 2048 > gpurun_out/chol_only.log 2>&1
python tools/acq_only.py 2000 524288 > gpurun_out/acq_only.log 2>&1
NCU="ncu --set full --import-source on --clock-control none -f"
$NCU -k regex:tile_kernel --launch-skip 15 --launch-count 2 -o gpurun_out/ncu_tile_potrf python tools/chol_only.py 592 2048 1 > gpurun_out/ncu1.log 2>&1
$NCU -k regex:tile_kernel --launch-skip 52 --launch-count 2 -o gpurun_out/ncu_tile_trtri python tools/chol_only.py 592 2048 1 > gpurun_out/ncu2.log 2>&1
$NCU -k "regex:gram_kernel|diag_kernel" -c 3 -o gpurun_out/ncu_gram_diag python tools/chol_only.py 592 2048 1 > gpurun_out/ncu3.log 2>&1
$NCU -k regex:acq_reduce -c 1 -o gpurun_out/ncu_acq python tools/acq_only.py 2000 262144 > gpurun_out/ncu4.log 2>&1
cat gpurun_out/chol_only.log gpurun_out/acq_only.log
