"""The few reference types the plugin API exchanges.  When the reference packages are importable (a GTSfM checkout
with gtsam on PYTHONPATH) their own classes are used, so the plugins are drop-ins; otherwise minimal stand-ins with
the same accessors keep the package usable (and testable) on a machine that only has this repository.

  Keypoints     gtsfm/common/keypoints.py:15-43        coordinates (N,2), optional scales / responses
  Cal3Bundler   gtsam (un-vendored, gtsam==4.2)        fx(), k1(), k2(), px(), py(), K()
  Rot3, Unit3   gtsam                                  Rot3(R).matrix(), Unit3(v).point3()
  MatcherBase / VerifierBase come from gtsfm.ui's GTSFMProcess registry when available.
"""
from typing import Optional

import numpy as np

try:  # pragma: no cover - exercised only inside a GTSfM environment
    from gtsfm.common.keypoints import Keypoints  # type: ignore
except Exception:  # noqa: BLE001
    class Keypoints:  # type: ignore[no-redef]
        """Stand-in for gtsfm.common.keypoints.Keypoints."""

        def __init__(self, coordinates: np.ndarray, scales: Optional[np.ndarray] = None, responses: Optional[np.ndarray] = None):
            self.coordinates = coordinates
            self.scales = scales
            self.responses = responses

        def __len__(self) -> int:
            return int(self.coordinates.shape[0])

        def extract_indices(self, indices: np.ndarray) -> "Keypoints":
            if indices.size == 0:
                return Keypoints(coordinates=np.zeros((0, 2)))
            return Keypoints(self.coordinates[indices], None if self.scales is None else self.scales[indices],
                             None if self.responses is None else self.responses[indices])

try:  # pragma: no cover
    from gtsam import Cal3Bundler, Rot3, Unit3  # type: ignore
    HAVE_GTSAM = True
except Exception:  # noqa: BLE001
    HAVE_GTSAM = False

    class Cal3Bundler:  # type: ignore[no-redef]
        def __init__(self, fx: float = 1.0, k1: float = 0.0, k2: float = 0.0, u0: float = 0.0, v0: float = 0.0):
            self._v = (float(fx), float(k1), float(k2), float(u0), float(v0))

        def fx(self): return self._v[0]
        def k1(self): return self._v[1]
        def k2(self): return self._v[2]
        def px(self): return self._v[3]
        def py(self): return self._v[4]

        def K(self):
            f, _, _, u0, v0 = self._v
            return np.array([[f, 0, u0], [0, f, v0], [0, 0, 1.0]])

    class Rot3:  # type: ignore[no-redef]
        def __init__(self, R=None):
            self._R = np.eye(3) if R is None else np.asarray(R, float).reshape(3, 3)

        def matrix(self): return self._R

    class Unit3:  # type: ignore[no-redef]
        def __init__(self, v=None):
            v = np.array([1.0, 0, 0]) if v is None else np.asarray(v, float).reshape(3)
            self._v = v / np.linalg.norm(v)

        def point3(self): return self._v

try:  # pragma: no cover
    from gtsfm.ui.gtsfm_process import GTSFMProcess, UiMetadata  # type: ignore
except Exception:  # noqa: BLE001
    from dataclasses import dataclass
    from typing import Tuple

    @dataclass(frozen=True)
    class UiMetadata:  # type: ignore[no-redef]
        display_name: str
        input_products: Tuple[str, ...]
        output_products: Tuple[str, ...]
        parent_plate: str = ""

    class GTSFMProcess:  # type: ignore[no-redef]
        """Stand-in for gtsfm.ui.gtsfm_process.GTSFMProcess (the registry metaclass is not needed outside GTSfM)."""


def cal_to_tuple(cal) -> tuple:
    """(f, k1, k2, u0, v0) of a Cal3Bundler-like object or a 5-sequence."""
    if hasattr(cal, "fx"):
        return (float(cal.fx()), float(cal.k1()), float(cal.k2()), float(cal.px()), float(cal.py()))
    c = tuple(float(x) for x in cal)
    if len(c) != 5:
        raise ValueError("calibration must be a Cal3Bundler or (f, k1, k2, u0, v0)")
    return c
