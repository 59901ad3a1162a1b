import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/ska-sdp-func_b200')
from tests import synth
from tests.helpers import check_parity
for name,c,eps in [("c1_f32", synth.reference_test_case(256, single=True),1e-5),
                   ("c2s", synth.uniform_box(250,256,1024,single=True),1e-5),
                   ("c5s", synth.wide_field(400,64,1024,single=True),1e-4),
                   ("c4s", synth.ska_low_snapshot(64,32,1024,single=True,r_max=4e3),1e-5)]:
    try: print(name, {k: float(np.float32(v)) for k,v in check_parity(c,eps,True).items()})
    except AssertionError as e: print(name,"FAIL",e)
