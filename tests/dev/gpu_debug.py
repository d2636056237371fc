"""Diagnostic run on a GPU box: CUDA engine vs the CPU oracle, layer by layer."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from lpcnet_b200 import blob, engine
from oracle import orcpy

def main():
    A = blob.random_model()
    b = blob.write_blob(A)
    O = orcpy.OrcLib()
    om = orcpy.OrcModel(O, b, 2, 1.0, None)
    print("devices", engine.device_count(), "nblocks", om.nblocks_a, om.nblocks_b, flush=True)

    # 1. frame network
    B, F = 5, 7
    feats = np.stack([blob.synthetic_features(F, s) for s in range(B)])
    bt = engine.Batch(b, B)
    ga, gb, lpc = bt.frame_network(feats)
    for s in range(B):
        st = orcpy.OrcStream(om)
        for f in range(F):
            oga, ogb, olpc = st.frame_network(feats[s, f])
            for nm, x, y in (("ga", ga[s, f], oga), ("gb", gb[s, f], ogb), ("lpc", lpc[s, f], olpc)):
                if not np.array_equal(x.view(np.uint32), y.view(np.uint32)):
                    print(f"FRAME MISMATCH s={s} f={f} {nm} maxabs={np.abs(x-y).max():.3e} nbad={(x!=y).sum()}")
    print("frame network checked", flush=True)
    bt.close()

    # 2. synthesis with trace
    for tile in (8, 16, 32, 64):
        for forced_mode in (False, True):
            B, F = 10, 5
            feats = np.stack([blob.synthetic_features(F, s) for s in range(B)])
            ref = []
            for s in range(B):
                st = orcpy.OrcStream(om, blob.stream_seed(s))
                ref.append(st.run(feats[s], trace=True, logits=True))
            forced = np.stack([r[0] for r in ref]) if forced_mode else None
            if forced_mode:
                ref = []
                for s in range(B):
                    st = orcpy.OrcStream(om, blob.stream_seed(s))
                    ref.append(st.run(feats[s], forced=forced[s], trace=True, logits=True))
            bt = engine.Batch(b, B)
            bt.set_tile(tile)
            for s in range(B):
                bt.set_seed(s, blob.stream_seed(s))
            t0 = time.time()
            pcm, tr, lg = bt.synthesize(feats, forced=forced, trace=True, logits=True)
            dt = time.time() - t0
            nbad = 0
            for s in range(B):
                rp, rt, rl = ref[s]
                ok = np.array_equal(pcm[s], rp) and np.array_equal(tr[s], rt)
                lgd = np.abs(lg[s] - rl).max()
                if not ok or lgd != 0:
                    nbad += 1
                    if nbad <= 3:
                        bad = np.nonzero(tr[s] != rt)[0]
                        i = bad[0] if len(bad) else -1
                        print(f"  tile={tile} forced={forced_mode} s={s}: pcm_eq={np.array_equal(pcm[s], rp)} first trace diff at {i}; logits maxdiff {lgd:.3e}")
                        if i >= 0:
                            print("    got", tr[s][i], "\n    ref", rt[i])
            print(f"tile={tile} forced={forced_mode}: {B-nbad}/{B} streams bit-exact ({dt:.2f}s)", flush=True)
            # state round trip
            st0 = bt.get_state(0)
            bt.close()

    # 3. timing
    for B, tile in ((1024, 8), (1024, 16), (4096, 32), (8192, 64)):
        F = 20
        feats = np.stack([blob.synthetic_features(F, s % 16) for s in range(B)])
        bt = engine.Batch(b, B)
        bt.set_tile(tile)
        bt.synthesize(feats[:, :4])
        t0 = time.time()
        pcm, _, _ = bt.synthesize(feats)
        dt = time.time() - t0
        print(f"B={B} tile={tile}: {F} frames in {dt*1e3:.1f} ms -> {B*F*160/dt:.3e} samples/s ({dt/(F*160)*1e6:.2f} us/step)", flush=True)
        bt.close()

if __name__ == "__main__":
    main()
