"""Quick GPU iteration check: parity on every tile size, then sample-kernel timing."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from lpcnet_b200 import blob, engine
from oracle import orcpy
if os.environ.get("LPCNET_B200_LIB"):       # A/B runs against another build of the library
    engine.LIB_PATH = os.environ["LPCNET_B200_LIB"]

def main():
    b = blob.write_blob(blob.random_model())
    if os.environ.get("QUICK_NO_PARITY"): return main_timing(b)
    om = orcpy.OrcModel(orcpy.OrcLib(), b, 2, 1.0, None)
    B, F = 19, 4
    feats = np.stack([blob.synthetic_features(F, s) for s in range(B)])
    seeds = [blob.stream_seed(s) for s in range(B)]
    free = [orcpy.OrcStream(om, seeds[s]).run(feats[s], trace=True, logits=True) for s in range(B)]
    forced = np.stack([(r[0].astype(np.int32) // 2).astype(np.int16) for r in free])
    frc = [orcpy.OrcStream(om, seeds[s]).run(feats[s], forced=forced[s], trace=True, logits=True) for s in range(B)]
    allok = True
    bt = engine.Batch(b, B)
    print("smem per CTA:", {t: bt.L.lpcnet_batch_nblocks(bt.h, t) for t in (8, 16, 32)}, "chunks A", bt.L.lpcnet_batch_nblocks(bt.h, 4), flush=True)
    tiles = [t for t in (8, 16, 32) if bt.L.lpcnet_batch_nblocks(bt.h, t) <= 232448]
    bt.close()
    for tile in tiles:
        for fm, ref in ((None, free), (forced, frc)):
            bt = engine.Batch(b, B); bt.set_tile(tile)
            for s in range(B): bt.set_seed(s, seeds[s])
            pcm, tr, lg = bt.synthesize(feats, forced=fm, trace=True, logits=True)
            ok = sum(np.array_equal(pcm[s], ref[s][0]) and np.array_equal(tr[s], ref[s][1]) and
                     np.array_equal(lg[s].view(np.uint32), ref[s][2].view(np.uint32)) for s in range(B))
            # the non-debug kernel too
            bt.reset()
            for s in range(B): bt.set_seed(s, seeds[s])
            pcm2, _, _ = bt.synthesize(feats, forced=fm)
            ok2 = sum(np.array_equal(pcm2[s], ref[s][0]) for s in range(B))
            print(f"tile={tile} forced={fm is not None}: debug {ok}/{B}  release {ok2}/{B}", flush=True)
            if ok != B or ok2 != B:
                allok = False
                for s in range(B):
                    bad = np.nonzero(tr[s] != ref[s][1])[0]
                    if len(bad):
                        print("  stream", s, "first diff step", bad[0], tr[s][bad[0]], ref[s][1][bad[0]]); break
            flags = (bt.rcp_arith, bt.fixed_slots)
            bt.close()
    print("PARITY", "OK" if allok else "FAILED", " rcp_arith, fixed_slots =", flags, flush=True)
    if os.environ.get("QUICK_NO_TIMING"): return
    main_timing(b)

def main_timing(b):
    cfgs = [(1024, 8), (2048, 16), (2368, 16), (4736, 32)]
    if len(sys.argv) > 1:
        cfgs = [tuple(int(x) for x in a.split(":")) for a in sys.argv[1:]]
    for Bn, tile in cfgs:
        F = 34
        feats = np.stack([blob.synthetic_features(F, s % 16) for s in range(Bn)])
        bt = engine.Batch(b, Bn); bt.set_tile(tile)
        bt.synthesize(feats[:, :4])
        bt.set_profiling(True); bt.read_profile()
        bt.synthesize(feats[:, :32])
        sm, sn, fm_, fn = bt.read_profile()
        steps = 32 * 160
        print(f"B={Bn} tile={tile}: sample kernel {sm:.1f} ms ({sm/steps*1e3:.2f} us/step, {Bn*steps/sm*1e3:.3e} samples/s), frame net {fm_:.2f} ms", flush=True)
        bt.close()

if __name__ == "__main__":
    main()
