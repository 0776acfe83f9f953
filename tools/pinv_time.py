import time, os, sys
sys.path.insert(0, os.getcwd())
import wenoext_b200 as wx
os.environ["WENOEXT_SETUP_TIMING"] = "1"
for gp in (False, True, True):
    t = time.time(); c = wx.BoxCase((20, 20, 20), 2, tetSplit=6, useClassCache=False, gpuPinv=gp); print("tets 20^3x6 gpuPinv", gp, round(time.time() - t, 2), "s", c.nCells, flush=True)
