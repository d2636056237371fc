"""Per-frame comparison of the CUDA PLC with the oracle, outputs and frame-rate state (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from lpcnet_b200 import blob, engine
from oracle import orcpy
g = np.load("tests/golden/plc.npz")
A = blob.random_model(); A.update(blob.random_plc_model()); b = blob.write_blob(A)
opt = int(sys.argv[1]) if len(sys.argv) > 1 else 1
om = orcpy.OrcModel(orcpy.OrcLib(), b, 0, 1.0, None)
for s in range(1):
    x, lost, want = g[f"in_{opt}_{s}"], g[f"lost_{opt}_{s}"], g[f"out_{opt}_{s}"]
    pb = engine.PLCBatch(b, 1, opt); pb.set_seed(0, blob.stream_seed(s))
    op = orcpy.OrcPLC(om, opt, blob.stream_seed(s))
    lc = 0
    for f in range(2):
        cls = ("L0" if lc == 0 else "L1") if lost[f] else ("G0" if lc == 0 else "FG")
        lc = lc + 1 if lost[f] else 0
        fr = x[f*160:(f+1)*160]
        got = pb.process(fr[None], lost[None, f:f+1])[0]
        w = op.run(fr, lost[f:f+1])
        pf, pg1, pg2, pdc = pb.peek(0); of, og1, og2, odc = op.peek()
        bg, en = pb.peek_frame(0)
        ob = orcpy.burg_cepstral_analysis(om.L, fr.astype(np.float32))
        if f == 0: print("gpu burg", bg, "\norc burg", ob)
        print("   burg", np.array_equal(bg.view(np.uint32), ob.view(np.uint32)), float(np.abs(bg - ob).max()), "enc[:4]", en[:4], en[18:20])
        print(f, cls, "out", np.array_equal(got[:80], w[:80]), np.array_equal(got[80:], w[80:]),
              "feat", np.array_equal(pf.view(np.uint32), of.view(np.uint32)), float(np.abs(pf - of).max()),
              "gru1", np.array_equal(pg1.view(np.uint32), og1.view(np.uint32)), float(np.abs(pg1 - og1).max()),
              "gru2", float(np.abs(pg2 - og2).max()), "dc", pdc, odc)
