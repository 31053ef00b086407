import numpy as np, sys
import of_dis_b200 as gpu
from oracle import bindings as ob
from of_dis_b200.synth import make_pair
from tests.helpers import base_orc, to_dis
port = ob.Port()
h0, w0 = int(sys.argv[1]), int(sys.argv[2])
A, B, _ = make_pair(h0, w0, seed=5)
q = base_orc(1, 1, sc_f=0, sc_l=0, tv_solverit=3)
a, b = port.pyramid_level(A, 0, 0, 8)[0], port.pyramid_level(B, 0, 0, 8)[0]
fl = (np.random.default_rng(1).standard_normal((h0, w0, 2)) * 0.5).astype(np.float32)
got = gpu.varref_level(to_dis(gpu, q), a, b, w0, h0, 0, 0, fl)
print("done", got.shape)
