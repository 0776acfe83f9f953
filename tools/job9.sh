mkdir -p gpurun_out
WENOEXT_SETUP_TIMING=1 timeout 900 python -m pytest tests/test_gpu_pinv.py -x -q > gpurun_out/ls_test.log 2>&1; echo "rc=$?" >> gpurun_out/ls_test.log
tail -25 gpurun_out/ls_test.log
WENOEXT_SETUP_TIMING=1 python - > gpurun_out/ls_timing.log 2>&1 <<'PY'
import time, wenoext_b200 as wx
for kw in (dict(gpuPinv=False), dict(gpuPinv=True, device=0), dict(gpuPinv=True, device=0)):
    t = time.time(); c = wx.BoxCase((40, 40, 40), 2, useClassCache=False, **kw); print(kw, "generic setup 40^3: %.2f s" % (time.time() - t), flush=True)
PY
tail -30 gpurun_out/ls_timing.log
