"""Per-phase cycle breakdown of the sample kernel (development build with -DLPCB_PHASE_CLOCKS).
   python -m lpcnet_b200.build --phase-clocks ; python tools/phase_clocks.py <streams> <tile> [frames]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lpcnet_b200 import blob, engine
engine.LIB_PATH = os.path.join(engine.HERE, "liblpcnet_b200_clk.so")
NAMES = ["outside step loop", "A (own work)", "A barrier wait", "B1 (own work)", "B1 barrier wait", "B2 GRU-B gates",
         "T round 1", "T round 2", "T signal update", "end barrier wait"]
B, tile = int(sys.argv[1]), int(sys.argv[2])
F = int(sys.argv[3]) if len(sys.argv) > 3 else 8
b = blob.write_blob(blob.random_model())
feats = np.stack([blob.synthetic_features(F + 2, s % 16) for s in range(B)])
bt = engine.Batch(b, B); bt.set_tile(tile)
bt.synthesize(feats[:, :2])
out = (ctypes.c_ulonglong * 16)()
L = engine.load_library()
L.lpcnet_b200_read_phase_clocks(out)
bt.synthesize(feats[:, 2:])
L.lpcnet_b200_read_phase_clocks(out)
steps = F * 160
tot = sum(out[1:10])
print(f"B={B} tile={tile}: {tot / steps:.0f} cycles/step (thread 0 of CTA 0)")
for k in range(10):
    print(f"  {NAMES[k]:22s} {out[k] / steps:8.0f}  {100.0 * out[k] / max(tot, 1):5.1f}%")
