"""A/B helper: run a fixed set of configurations with the library named by DISB200_LIB and save the flows.
   python scratch/ab_identity.py out.npz          (then compare two files with --cmp a.npz b.npz)"""
import sys
import numpy as np

if sys.argv[1] == "--cmp":
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    ok = True
    for k in a.files:
        same = np.array_equal(a[k], b[k])
        ok &= same
        print(k, "identical" if same else "DIFFERENT max %g" % np.abs(a[k] - b[k]).max(), "mean |flow| %.3f" % np.abs(a[k]).mean())
    sys.exit(0 if ok else 1)

import of_dis_b200 as dis
from of_dis_b200.synth import make_pair

out = {}
cases = [("op4", 1, 1, 4, 512, 218, {}), ("op3", 1, 1, 3, 1024, 436, {}), ("op4_rgb", 1, 3, 4, 320, 200, {}), ("op4_de", 2, 1, 4, 640, 360, {}),
         ("op2_rgb", 1, 3, 2, 1024, 436, {}), ("op2_rgb_iter64", 1, 3, 2, 512, 218, dict(max_iter=64, min_iter=64)),
         ("psz16_iter100", 1, 1, 2, 512, 218, dict(patchsz=16, max_iter=100, min_iter=40)), ("psz4_iter50", 1, 1, 2, 512, 218, dict(patchsz=4, max_iter=50, min_iter=50)),
         ("op4_minlt", 1, 1, 4, 512, 218, dict(min_iter=33, max_iter=90)), ("op4_fb_l1", 1, 1, 4, 320, 200, dict(usefbcon=1, costfct=1))]
for name, mode, ch, op, w0, h0, kw in cases:
    n = 3
    pairs = [make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=70 + s)[:2] for s in range(n)]
    A, B = np.stack([a for a, _ in pairs]), np.stack([b for _, b in pairs])
    p = dis.operating_point(op, w0, mode, ch).copy(verbosity=0, **kw)
    with dis.Handle(p, w0, h0, n) as hd:
        out[name] = hd.run_host(A, B)
np.savez(sys.argv[1], **out)
print("saved", sys.argv[1], dis.LIB_PATH)
