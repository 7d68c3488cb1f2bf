"""Retriever step: similarity matrix + top-K pair selection.

Golden: tests/golden/retriever.npz, recorded from the unmodified reference NetVLADRetriever on CPU torch
(oracle/gen_golden_retriever.py).  The CPU tests pin the numpy oracle against it; the GPU tests compare the device path with
both.  Scores are float32 dot products whose summation order is BLAS-defined in the reference, so the matrix is compared to
1e-5 and the pair lists exactly (the scores that decide a fixture row's top K are at least 2e-6 apart, checked below).
"""
import os

import numpy as np
import pytest

TOL = 1e-5


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "retriever.npz"))


def case(g, k):
    p = f"r{k}_"
    ms = float(g[p + "min_score"])
    return g[p + "g"], int(g[p + "num_matched"]), (None if np.isnan(ms) else ms), int(g[p + "blocksize"]), g[p + "sim"], g[p + "pairs"]


@pytest.mark.parametrize("k", range(6))
def test_oracle_reproduces_reference_retriever(gold, k):
    from oracle import oracle as O
    G, num_matched, min_score, blocksize, sim_ref, pairs_ref = case(gold, k)
    sim = O.retriever_similarity(G, blocksize)
    np.testing.assert_allclose(sim, sim_ref, atol=TOL, rtol=0)
    assert (sim_ref[np.tril_indices(len(G), -blocksize)] == 0).all()          # far below the block diagonal: never computed
    # the scores that decide a row's top K are well separated, so the pair list is unambiguous
    n = len(G)
    for i in range(n - 2):
        top = np.sort(sim_ref[i, i + 1:])[::-1][:num_matched + 1]
        assert top.size < 2 or (-np.diff(top)).min() > 2e-6
    pairs = O.pairs_from_score_matrix(sim_ref, num_matched, min_score)
    np.testing.assert_array_equal(np.array(pairs, np.int32).reshape(-1, 2), pairs_ref)
    pairs2 = O.pairs_from_score_matrix(sim, num_matched, min_score)
    np.testing.assert_array_equal(np.array(pairs2, np.int32).reshape(-1, 2), pairs_ref)


@pytest.mark.gpu
@pytest.mark.parametrize("k", range(6))
def test_device_retriever_equals_reference(gold, k):
    from gtsfm_b200.retriever.netvlad_retriever import NetVLADRetriever
    G, num_matched, min_score, blocksize, sim_ref, pairs_ref = case(gold, k)
    r = NetVLADRetriever(num_matched=num_matched, min_score=min_score, blocksize=blocksize)
    descs = [G[i] for i in range(len(G))]
    names = [f"{i}.jpg" for i in range(len(G))]
    sim = r.compute_similarity_matrix(descs)
    assert sim.dtype == np.float32 and sim.shape == sim_ref.shape
    np.testing.assert_allclose(sim, sim_ref, atol=TOL, rtol=0)
    assert ((sim == 0) == (sim_ref == 0)).all()                               # same computed / never-computed pattern
    pairs = r.get_image_pairs(descs, names)
    np.testing.assert_array_equal(np.array(pairs, np.int32).reshape(-1, 2), pairs_ref)
    assert all(i < j for i, j in pairs)
    pairs_b = r.compute_pairs_from_similarity_matrix(sim_ref, names)          # caller-supplied matrix (the reference's, verbatim)
    np.testing.assert_array_equal(np.array(pairs_b, np.int32).reshape(-1, 2), pairs_ref)


@pytest.mark.gpu
def test_device_retriever_larger_case_vs_oracle_and_artifacts(tmp_path):
    """601 images x 4,096-d (NetVLAD's dimension), ragged tile edges, ties; plus the text artefacts of plots_output_dir."""
    from oracle import oracle as O
    from gtsfm_b200.retriever.netvlad_retriever import NetVLADRetriever
    rng = np.random.default_rng(3)
    n, dim = 601, 4096
    G = rng.standard_normal((n, dim)).astype(np.float32)
    G /= np.linalg.norm(G, axis=1, keepdims=True)
    G[100] = G[7]                                                             # exact duplicates: equal scores in many rows
    G[300] = G[7]
    r = NetVLADRetriever(num_matched=20, min_score=0.0, blocksize=50)
    names = [f"im{i}.png" for i in range(n)]
    sim = r.compute_similarity_matrix(list(G))
    ref = O.retriever_similarity(G, 50)
    np.testing.assert_allclose(sim, ref, atol=TOL, rtol=0)
    pairs = r.get_image_pairs(list(G), names, plots_output_dir=tmp_path)
    want = O.pairs_from_score_matrix(sim, 20, 0.0)                            # same scores -> the tie rule is what is tested
    assert [(int(i), int(j)) for i, j in pairs] == want
    row7 = [j for i, j in want if i == 7]
    assert row7[:2] == [100, 300]                                             # duplicates of image 7: score 1, lower index first
    assert (tmp_path / "netvlad_similarity_matrix.txt").exists()
    lines = (tmp_path / "netvlad_named_pairs.txt").read_text().splitlines()
    assert len(lines) == len(pairs) and lines[0].split()[1:] == [names[pairs[0][0]], names[pairs[0][1]]]


@pytest.mark.gpu
def test_device_retriever_errors():
    from gtsfm_b200.retriever.netvlad_retriever import NetVLADRetriever
    r = NetVLADRetriever(num_matched=2)
    with pytest.raises(ValueError):
        r.get_image_pairs(None, ["a", "b"])
    with pytest.raises(RuntimeError):
        r.compute_similarity_matrix([np.zeros(4, np.float32)] * 10001)
    # one image: no pairs
    assert r.get_image_pairs([np.ones(8, np.float32)], ["a"]) == []


@pytest.mark.gpu
@pytest.mark.parametrize("n,dim,k,blocksize", [(77, 100, 4, 50), (65, 7, 70, 3), (2, 513, 1, 50), (129, 130, 5, 64)])
def test_device_retriever_odd_shapes_vs_oracle(n, dim, k, blocksize):
    """Dimensions that are not multiples of the 16-wide K chunk, image counts around the 64-wide tile, tiny blocks."""
    from oracle import oracle as O
    from gtsfm_b200.retriever.netvlad_retriever import NetVLADRetriever
    rng = np.random.default_rng(n * 1000 + dim)
    G = rng.standard_normal((n, dim)).astype(np.float32)
    G /= np.linalg.norm(G, axis=1, keepdims=True)
    r = NetVLADRetriever(num_matched=k, min_score=-1.0, blocksize=blocksize)
    sim = r.compute_similarity_matrix(list(G))
    ref = O.retriever_similarity(G, blocksize)
    np.testing.assert_allclose(sim, ref, atol=TOL, rtol=0)
    assert ((sim == 0) == (ref == 0)).all() or np.abs(ref[(sim == 0) != (ref == 0)]).max() < TOL
    pairs = r.get_image_pairs(list(G), [str(i) for i in range(n)])
    assert [(int(i), int(j)) for i, j in pairs] == O.pairs_from_score_matrix(sim, k, -1.0)
