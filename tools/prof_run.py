"""Small fixed workload for ncu: python tools/prof_run.py <streams> <tile> [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lpcnet_b200 import blob, engine
B, tile = int(sys.argv[1]), int(sys.argv[2])
F = int(sys.argv[3]) if len(sys.argv) > 3 else 6
b = blob.write_blob(blob.random_model())
feats = np.stack([blob.synthetic_features(F, s % 16) for s in range(B)])
bt = engine.Batch(b, B); bt.set_tile(tile)
bt.synthesize(feats)     # launch 1 (warm: includes 2 inactive frames)
bt.synthesize(feats)     # launch 2: all frames active
print("ok", bt.launch_count)
