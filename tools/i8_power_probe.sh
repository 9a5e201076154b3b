#!/bin/bash
# SM clock / power while the INT8 predictive sweep runs back to back (is it power-capped?)
#   tools/i8_power_probe.sh [GPX_I8_DEBUG bits]
export GPX_I8_DEBUG=${1:-0}
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap --format=csv,noheader -lms 100 > /tmp/i8_smi_$GPX_I8_DEBUG.log &
SMI=$!
python - <<'PY'
import os, sys, time
sys.path.insert(0, os.getcwd())
import torch
from gpax_b200 import engine as E, synthetic as syn
wl = syn.config2(n=2048, S=74, grid=100)
dev = torch.device("cuda:0")
post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
xn = E._dev(wl.x_new, dev)
ws = E.predict_i8_ws_alloc(74, 2048, wl.d, dev)
mu, var = E.predict_i8(post.hpd, post.x, xn, post.W, post.alpha, post.hpd.jitter, ws=ws)
torch.cuda.synchronize()
time.sleep(1.0)
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(60):
    E.predict_i8(post.hpd, post.x, xn, post.W, post.alpha, post.hpd.jitter, mu=mu, var=var, ws=ws, prepare=False)
e1.record(); torch.cuda.synchronize()
print("dbg", os.environ["GPX_I8_DEBUG"], "ms/sweep", e0.elapsed_time(e1) / 60)
PY
kill $SMI
awk -F, '{gsub(/ W/,"",$2); if ($2+0 > 400) print}' /tmp/i8_smi_$GPX_I8_DEBUG.log | sort | uniq -c | sort -rn | head -8
