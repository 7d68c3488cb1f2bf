"""The device-resident result records (csrc/result_pack.cu): headers, prefix-sum offsets and the compacted
v_corr_idxs payload of a verify batch equal, byte for byte, the host mirror built from the fetched results; the single-GPU
exchange path (pack -> torch tensor -> pinned host) returns the same dictionaries as the per-batch result."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_device_pack_equals_host_mirror_and_exchange_roundtrip():
    import torch
    import gtsfm_b200
    from gtsfm_b200 import _native, synthetic
    from gtsfm_b200.two_view_batch import ResultExchange, TwoViewBatchResult, block_bytes, block_layout, pack_block
    scene = synthetic.make_scene(n_images=7, n_kps=2048, seed=31)
    descs, kps = list(scene.descriptors), list(scene.keypoints)
    descs[5] = descs[5][:4]; kps[5] = kps[5][:4]                     # pairs with image 5: fewer than 6 matches -> status 1
    ctx = gtsfm_b200.Context(0)
    for i in range(7):
        ctx.upload_image(i, descs[i], kps[i], scene.cal)
    pairs = scene.pairs
    P = len(pairs)
    prm = _native.VerifyParams()
    prm.threshold_px, prm.n_hypotheses, prm.lo_iters, prm.seed = 4.0, 256, 20, 5
    m, off = ctx.match_batch(pairs, 0.8)
    out = ctx.verify_batch(pairs, prm, total=len(m))
    res = TwoViewBatchResult(pairs, off, m, out["mask"], out["R"], out["t"], out["n_inliers"], out["status"])
    assert (out["status"] == 1).any() and (out["status"] == 0).any()
    ids = np.arange(P, dtype=np.int64)[::-1].copy() + 100             # arbitrary global ids
    hdr, payload = pack_block(res, ids)
    # ---- device pack into a torch tensor
    rows = ctx.result_pack(P, ids)
    assert rows == len(payload) == int(hdr["n_inliers"].sum())
    p_off, size = block_layout(P + 3, rows + 5)                       # padded like a smaller rank in an all-gather
    block = torch.full((size,), 0, dtype=torch.uint8, device="cuda")
    ctx.result_pack_copy(block.data_ptr(), size, p_off)
    ctx.sync()
    got = block.cpu().numpy()
    want = block_bytes(hdr, payload, P + 3, rows + 5)
    np.testing.assert_array_equal(got[: 128 * P], want[: 128 * P])
    np.testing.assert_array_equal(got[p_off: p_off + 8 * rows], want[p_off: p_off + 8 * rows])
    with pytest.raises(gtsfm_b200.GtsfmB200Error):                    # destination too small
        ctx.result_pack_copy(block.data_ptr(), p_off + 8 * rows - 8, p_off)
    # ---- the exchange on one process: pack -> pinned host -> dictionaries
    ex = ResultExchange(None)
    ctx.verify_batch(pairs, prm, total=len(m), fetch=False)            # results stay on the device
    table = ex.finish(ex.start_device(ctx, P, np.arange(P, dtype=np.int64)), P)
    Rd, td, vd, rd = table.two_view_dicts(pairs)
    Rw, tw, vw, rw = res.two_view_dicts()
    for key in Rw:
        assert (Rd[key] is None) == (Rw[key] is None)
        np.testing.assert_array_equal(vd[key], vw[key])
        assert abs(rd[key] - rw[key]) < 1e-15
        if Rw[key] is not None:
            np.testing.assert_array_equal(Rd[key].matrix(), Rw[key].matrix())
            np.testing.assert_allclose(td[key].point3(), tw[key].point3(), atol=1e-15)
    ctx.close()
