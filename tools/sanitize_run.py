"""Small workloads for compute-sanitizer (memcheck / racecheck / synccheck): every batched entry point once, tiny sizes.
    compute-sanitizer --tool racecheck python tools/sanitize_run.py [synth|plc|codec|all]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lpcnet_b200 import blob, engine

what = sys.argv[1] if len(sys.argv) > 1 else "all"
A = blob.random_model()
A.update(blob.random_plc_model())
bl = blob.write_blob(A)
if what in ("synth", "all"):
    for tile, B in ((8, 11), (16, 19), (32, 35)):
        bt = engine.Batch(bl, B); bt.set_tile(tile)
        f = np.stack([blob.synthetic_features(3, s) for s in range(B)])
        bt.synthesize(f); bt.synthesize(f, trace=True, logits=True)
        bt.close()
        print("synth tile", tile, "ok", flush=True)
if what in ("plc", "all"):
    for opt in (5, 4):
        B, F = 24, 6
        pcm = np.stack([blob.synthetic_pcm(F, 100 + s) for s in range(B)])
        lost = np.stack([blob.synthetic_loss_pattern(F, 7 + s, 0.5, 0.3) for s in range(B)])
        pb = engine.PLCBatch(bl, B, opt)
        pb.process_inplace(pcm, lost)
        pb.close()
        print("plc options", opt, "ok", flush=True)
if what in ("codec", "all"):
    B = 9
    fb = engine.FeatureBatch(B)
    fb.set_codebooks(blob.synthetic_codebooks())
    pcm = np.stack([blob.synthetic_pcm(8, 100 + s) for s in range(B)])
    pk = fb.encode(pcm)
    fb.compute(pcm)
    fb.close()
    bt = engine.Batch(bl, B); bt.set_codebooks(blob.synthetic_codebooks())
    bt.decode(np.asarray(pk[0] if isinstance(pk, tuple) else pk))
    bt.close()
    print("codec ok", flush=True)
