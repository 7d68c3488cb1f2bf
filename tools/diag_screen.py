"""First-light diagnostic of the tcgen05 screening kernel: compares its raw outputs with numpy for one pair."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gtsfm_b200

def run(n1, n2, dim, seed=0, hi=256):
    rng = np.random.default_rng(seed)
    d1 = rng.integers(0, hi, (n1, dim)).astype(np.float32)
    d2 = rng.integers(0, hi, (n2, dim)).astype(np.float32)
    scale = 1.0
    if dim == 128 and hi == 256:     # keep |x|^2 < 2^23
        d1 = np.floor(d1 * 0.6); d2 = np.floor(d2 * 0.6)
    c = gtsfm_b200.Context(0)
    c.clear(); c.upload_image(0, d1); c.upload_image(1, d2)
    m, off = c.match_batch(np.array([[0, 1]], np.int32), None)
    n1p = (n1 + 255) // 256 * 256; n2p = (n2 + 255) // 256 * 256
    nhb = n2p // 128
    v1 = c.debug_fetch(0, nhb * n1p, np.float32).reshape(nhb, n1p)
    v2 = c.debug_fetch(1, nhb * n1p, np.float32).reshape(nhb, n1p)
    g = c.debug_fetch(2, nhb * n1p, np.uint8).reshape(nhb, n1p)
    cr = c.debug_fetch(3, n2p * 128, np.float32).reshape(n2p, 128)
    nflag = c.debug_fetch(10, 1, np.int32)[0]
    D = (d1.astype(np.float64) ** 2).sum(1)[:, None] + (d2.astype(np.float64) ** 2).sum(1)[None, :] - 2 * d1.astype(np.float64) @ d2.astype(np.float64).T
    H = np.full((n1p, n2p), 2.0 ** 99)       # padding rows are huge
    H[:n1, :n2] = D / 2
    G = H.reshape(n1p, nhb, 8, 16).min(3)                    # group minima
    Gs = np.sort(G, axis=2)
    e_v1 = Gs[:, :, 0].T; e_v2 = Gs[:, :, 1].T; e_g = G.argmin(2).T
    ok1 = (v1[:, :n1] == e_v1[:, :n1]); ok2 = (v2[:, :n1] == e_v2[:, :n1]); okg = (g[:, :n1] == e_g[:, :n1])
    # colraw[t][lane] = min over q = lane mod 128
    Hc = H.reshape(n1p // 128, 128, n2p).min(0).T            # [n2p][128]
    okc = (cr[:n2] == Hc[:n2])
    print(f"n1={n1} n2={n2} dim={dim}: rp_v1 ok {ok1.mean():.4f}  rp_v2 ok {ok2.mean():.4f}  rp_g ok {okg.mean():.4f}  colraw ok {okc.mean():.4f}  flagged {nflag}  matches {m.shape[0]}")
    if ok1.mean() < 1:
        bad = np.argwhere(~ok1)[:8]
        for hb, i in bad:
            print("   v1 mismatch hb", hb, "row", i, "got", v1[hb, i], "want", e_v1[hb, i], "ratio", v1[hb, i] / max(e_v1[hb, i], 1e-9))
        print("   rows bad by (row%128<64?):", (~ok1)[:, : n1 // 128 * 128].reshape(nhb, -1, 128).mean((0, 1))[:8], "...")
    if okc.mean() < 1:
        bad = np.argwhere(~okc)[:8]
        for t, l in bad:
            print("   colraw mismatch col", t, "lane", l, "got", cr[t, l], "want", Hc[t, l])
    c.close()
    return ok1.all() and ok2.all() and okg.all() and okc.all()

if __name__ == "__main__":
    allok = True
    for (a, b, d, hi) in [(256, 256, 64, 4), (256, 256, 128, 4), (256, 256, 128, 256), (1000, 777, 128, 256), (5000, 5000, 128, 256), (300, 300, 64, 256)]:
        try:
            allok &= bool(run(a, b, d, hi=hi))
        except Exception as e:
            print("EXC", a, b, d, repr(e)); allok = False
    print("ALL OK" if allok else "FAILURES")
