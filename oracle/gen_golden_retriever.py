"""Golden vectors for the retriever step, recorded from the UNMODIFIED reference
NetVLADRetriever (gtsfm/retriever/netvlad_retriever.py) imported from /root/reference and run on the CPU (torch).
ORACLE-side test infrastructure; run in the build container:

    python oracle/gen_golden_retriever.py        -> tests/golden/retriever.npz

Modules the reference imports but this image lacks (matplotlib, gtsam, ...) are replaced by MagicMock placeholders in this
generator only; none of them takes part in the similarity / top-K code.  The NetVLAD network itself (weights are downloaded
by the reference) is not needed: the inputs are synthetic unit-norm global descriptors with a neighbourhood structure.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden_cacher import OUT, import_with_placeholders  # noqa: E402


def descriptors(rng, n, dim, walk=0.35):
    """Unit-norm descriptors along a random walk: neighbours in index are similar, like consecutive frames."""
    g = np.zeros((n, dim))
    v = rng.standard_normal(dim)
    for i in range(n):
        v = v + walk * np.linalg.norm(v) * rng.standard_normal(dim) / np.sqrt(dim) * 3
        g[i] = v / np.linalg.norm(v)
    return g.astype(np.float32)


def main():
    mod, stubbed = import_with_placeholders("gtsfm.retriever.netvlad_retriever")
    rng = np.random.default_rng(20260044)
    out = {}
    cases = [(12, 64, 2, 0.1, 50), (60, 128, 5, 0.1, 50), (130, 256, 7, 0.3, 50), (75, 96, 100, None, 20), (40, 32, 3, 0.95, 7),
             (40, 32, 3, 0.6, 7)]
    for k, (n, dim, num_matched, min_score, blocksize) in enumerate(cases):
        g = descriptors(rng, n, dim)
        r = mod.NetVLADRetriever(num_matched=num_matched, min_score=min_score, blocksize=blocksize)
        sim = r.compute_similarity_matrix([g[i] for i in range(n)])
        sim_np = sim.detach().cpu().numpy().copy()
        pairs = r.compute_pairs_from_similarity_matrix(sim=sim, image_fnames=[f"{i}.jpg" for i in range(n)], plots_output_dir=None)
        p = f"r{k}_"
        out.update({p + "g": g, p + "num_matched": np.array(num_matched), p + "min_score": np.array(np.nan if min_score is None else min_score),
                    p + "blocksize": np.array(blocksize), p + "sim": sim_np.astype(np.float32),
                    p + "pairs": np.array([(int(i), int(j)) for i, j in pairs], np.int32).reshape(-1, 2)})
        print(k, n, dim, num_matched, min_score, blocksize, "->", len(pairs), "pairs")
    out["n_cases"] = np.array(len(cases))
    np.savez_compressed(os.path.join(OUT, "retriever.npz"), **out)
    import torch
    with open(os.path.join(OUT, "PROVENANCE.txt"), "a") as fh:
        fh.write("retriever.npz: oracle/gen_golden_retriever.py, /root/reference (commit bc04d78b), unmodified NetVLADRetriever on CPU "
                 f"torch {torch.__version__}, placeholder modules {sorted(set(stubbed))}, numpy {np.__version__}\n")


if __name__ == "__main__":
    main()
