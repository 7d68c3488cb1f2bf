"""gtsfm_b200 -- B200-native (sm_100a) two-view front-end for GTSfM: TwoWayMatcher + Ransac verifier, batched.

Only the hot path of the reference is here; everything is served by libgtsfm_b200.so through the C ABI
in include/gtsfm_b200.h.  There is no CPU fallback: using the package without the built library or without a B200
raises gtsfm_b200._native.GtsfmB200Error.
"""
from gtsfm_b200._native import Context, GtsfmB200Error, VerifyParams, default_context, ransac_samples  # noqa: F401

__all__ = ["Context", "GtsfmB200Error", "VerifyParams", "default_context", "ransac_samples"]
