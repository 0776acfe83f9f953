import numpy as np, sys
sys.path.insert(0,'/root/repo')
import wenoext_b200 as wx
from tests import helpers as H
case = wx.BoxCase((8, 8, 8), 2, useClassCache=False)
vf = H.sin_field(case.C); phi = H.rotating_flux(case)
out = wx.WENOBase.instance(case).correction(0, 0, vf, phi)
print(out[:5])
