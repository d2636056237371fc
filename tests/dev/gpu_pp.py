"""Two 8-stream tiles per CTA (tile code 28) against 16-stream tiles: parity vs the oracle, then timing."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from lpcnet_b200 import blob, engine
from oracle import orcpy

b = blob.write_blob(blob.random_model())
om = orcpy.OrcModel(orcpy.OrcLib(), b, 2, 1.0, None)
B, F = 43, 5
feats = np.stack([blob.synthetic_features(F, s) for s in range(B)])
seeds = [blob.stream_seed(s) for s in range(B)]
ref = [orcpy.OrcStream(om, seeds[s]).run(feats[s])[0] for s in range(B)]
forced = np.stack([(r.astype(np.int32) // 2).astype(np.int16) for r in ref])
fref = [orcpy.OrcStream(om, seeds[s]).run(feats[s], forced=forced[s])[0] for s in range(B)]
for fm, rr in ((None, ref), (forced, fref)):
    bt = engine.Batch(b, B); bt.set_tile(28)
    for s in range(B): bt.set_seed(s, seeds[s])
    pcm, _, _ = bt.synthesize(feats, forced=fm)
    ok = sum(np.array_equal(pcm[s], rr[s]) for s in range(B))
    print(f"tile=28 forced={fm is not None}: {ok}/{B}", flush=True)
    bt.close()
for Bn, tile in [(int(x.split(':')[0]), int(x.split(':')[1])) for x in (sys.argv[1:] or ["2368:16", "2368:28", "4736:28", "9472:28", "9472:16"])]:
    Fn = 34
    f = np.stack([blob.synthetic_features(Fn, s % 16) for s in range(Bn)])
    bt = engine.Batch(b, Bn); bt.set_tile(tile)
    bt.synthesize(f[:, :4])
    bt.set_profiling(True); bt.read_profile()
    bt.synthesize(f[:, :32])
    sm, sn, fm_, fn = bt.read_profile()
    steps = 32 * 160
    print(f"B={Bn} tile={tile}: sample kernel {sm:.1f} ms, {Bn*steps/sm*1e3:.3e} samples/s", flush=True)
    bt.close()
