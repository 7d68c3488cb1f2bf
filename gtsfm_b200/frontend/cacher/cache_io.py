"""On-disk cache plumbing shared by the cachers: where the cache lives, the hash, and the .pbz2 entry format.  Host only.

Mirrors what the reference cachers rely on, so entries written by either side are read by the other:
  hash         sha1 over the raw buffer of one numpy array           gtsfm/utils/cache.py:20-22
  entry        pickle of the result object inside a bz2 stream        gtsfm/utils/io.py:610-630
  corrupt file logged, deleted, treated as a miss                     gtsfm/utils/io.py:615-620
  location     <GTSfM checkout>/cache/<namespace>/<key>.pbz2          matcher_cacher.py:24,42-44; two_view_estimator_cacher.py:25,45-47
"""
import bz2
import hashlib
import logging
import os
import pickle
from pathlib import Path
from typing import Any, Optional

import numpy as np

logger = logging.getLogger(__name__)


def cache_root() -> Path:
    """$GTSFM_CACHE_ROOT if set; else the `cache` folder of the importable GTSfM checkout (what the reference uses);
    else ./cache."""
    env = os.environ.get("GTSFM_CACHE_ROOT")
    if env:
        return Path(env)
    try:  # pragma: no cover - only inside a GTSfM environment
        import gtsfm  # type: ignore
        return Path(gtsfm.__file__).resolve().parent.parent / "cache"
    except Exception:  # noqa: BLE001
        return Path.cwd() / "cache"


def hash_numpy(array: np.ndarray) -> str:
    return hashlib.sha1(array).hexdigest()


def read_entry(path: Path) -> Optional[Any]:
    if not path.exists():
        return None
    try:
        with bz2.BZ2File(path, "rb") as f:
            return pickle.load(f)
    except Exception:  # noqa: BLE001
        logger.exception("Cache file %s was corrupted, removing it...", path)
        os.remove(path)
        return None


def write_entry(data: Any, path: Path) -> None:
    path.parent.mkdir(exist_ok=True, parents=True)
    with bz2.BZ2File(path, "wb") as f:
        pickle.dump(data, f)
