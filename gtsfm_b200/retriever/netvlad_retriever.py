"""NetVLADRetriever on the device: global-descriptor similarity + per-image top-K pair selection, the step that produces the
pair list the two-view path consumes.  Mirrors gtsfm/retriever/netvlad_retriever.py:34-228 (same
constructor arguments, method names and return conventions); the arithmetic runs in libgtsfm_b200.so (csrc/retrieve.cu).

Differences that are visible to a caller: `compute_similarity_matrix` returns a numpy float32 array instead of a
torch.Tensor (`compute_pairs_from_similarity_matrix` accepts either); score ties inside one row are broken by the lower image
index (torch.topk leaves the order of ties unspecified); the similarity-matrix JPEG plot needs matplotlib and is skipped
without it (the .txt artefacts are always written).
"""
import os
from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np

from gtsfm_b200 import _native

try:  # pragma: no cover - inside a GTSfM environment the class registers like the reference's (retriever_base.py:26-46)
    from gtsfm.retriever.retriever_base import ImageMatchingRegime, RetrieverBase  # type: ignore
except Exception:  # noqa: BLE001
    class RetrieverBase:  # type: ignore[no-redef]
        """Stand-in for gtsfm.retriever.retriever_base.RetrieverBase."""

        def __init__(self, matching_regime) -> None:
            self._matching_regime = matching_regime

    class ImageMatchingRegime:  # type: ignore[no-redef]
        RETRIEVAL = "retrieval"

MAX_NUM_IMAGES = 10000


class NetVLADRetriever(RetrieverBase):
    def __init__(self, num_matched: int, min_score: float = 0.1, blocksize: int = 50, device: int = 0) -> None:
        """
        Args:
            num_matched: number of potential matches to provide per query (the top K per image).
            min_score: minimum allowed similarity score to accept a match.
            blocksize: sub-block size of the reference's blocked similarity computation; only decides which entries of the
                returned matrix are filled (blocks on or above the block diagonal), never the pairs.
        """
        super().__init__(matching_regime=ImageMatchingRegime.RETRIEVAL)
        self._num_matched = num_matched
        self._blocksize = blocksize
        self._min_score = min_score
        self._device = device

    def __repr__(self) -> str:
        return f"""
        NetVLADRetriever:
            Num. frames matched: {self._num_matched}
            Block size: {self._blocksize}
            Minimum score: {self._min_score}
        """

    def get_image_pairs(self, global_descriptors: Optional[List[np.ndarray]], image_fnames: List[str],
                        plots_output_dir: Optional[Path] = None) -> List[Tuple[int, int]]:
        if global_descriptors is None:
            raise ValueError("Global descriptors need to be provided")                       # netvlad_retriever.py:70-71
        ctx = _native.default_context(self._device)
        g = self._stack(global_descriptors)
        if plots_output_dir:
            sim = ctx.retriever_similarity(g, self._blocksize, fetch=True)
            return self.compute_pairs_from_similarity_matrix(sim, image_fnames, plots_output_dir, _on_device=True)
        ctx.retriever_similarity(g, self._blocksize, fetch=False)                            # the matrix never leaves the device
        return self._pairs(ctx, None, len(image_fnames))

    @staticmethod
    def _stack(global_descriptors) -> np.ndarray:
        if len(global_descriptors) > MAX_NUM_IMAGES:
            raise RuntimeError("Cannot construct similarity matrix of this size.")           # netvlad_retriever.py:91-92
        return np.ascontiguousarray(np.array(global_descriptors), np.float32).reshape(len(global_descriptors), -1)

    def compute_similarity_matrix(self, global_descriptors: List[np.ndarray]) -> np.ndarray:
        """(num_images, num_images) float32; entries below the block diagonal are 0 as in the reference."""
        return _native.default_context(self._device).retriever_similarity(self._stack(global_descriptors), self._blocksize, fetch=True)

    def _pairs(self, ctx, sim: Optional[np.ndarray], num_images: int) -> List[Tuple[int, int]]:
        idx, _ = ctx.retriever_topk(sim, num_images, self._num_matched, self._min_score)
        rows, slots = np.nonzero(idx >= 0)                                                   # row-major, as np.where(valid)
        return [(i, j) for i, j in zip(rows, idx[rows, slots])]

    def compute_pairs_from_similarity_matrix(self, sim, image_fnames: List[str], plots_output_dir: Optional[Path] = None,
                                             _on_device: bool = False) -> List[Tuple[int, int]]:
        num_images = len(image_fnames)
        if hasattr(sim, "detach"):
            sim = sim.detach().cpu().numpy()
        sim = np.asarray(sim, np.float32)
        ctx = _native.default_context(self._device)
        pairs = self._pairs(ctx, None if _on_device else sim, num_images)
        if plots_output_dir:
            plots_output_dir = Path(plots_output_dir)
            os.makedirs(plots_output_dir, exist_ok=True)
            try:  # pragma: no cover - matplotlib is optional
                import matplotlib.pyplot as plt
                plt.imshow(np.triu(sim))
                plt.title("Image Similarity Matrix")
                plt.savefig(str(plots_output_dir / "netvlad_similarity_matrix.jpg"), dpi=500)
                plt.close("all")
            except ImportError:
                pass
            np.savetxt(fname=str(plots_output_dir / "netvlad_similarity_matrix.txt"), X=sim, fmt="%.2f", delimiter=",")
            with open(plots_output_dir / "netvlad_named_pairs.txt", "w") as fid:
                for i, j in pairs:
                    fid.write("%.4f %s %s\n" % (sim[i, j], image_fnames[i], image_fnames[j]))
        return pairs
