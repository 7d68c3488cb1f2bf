"""ctypes binding of libgtsfm_b200.so (the C ABI declared in include/gtsfm_b200.h).

The library is the product: there is no Python or CPU fallback.  Importing this module never touches the GPU; creating
a Context does, and raises GtsfmB200Error when the shared library or a B200 is missing.
"""
import ctypes
import os
import threading
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgtsfm_b200.so")

# every symbol include/gtsfm_b200.h declares
EXPORTS = (
    "gtsfm_ctx_create", "gtsfm_ctx_destroy", "gtsfm_last_error", "gtsfm_ctx_set_workspace", "gtsfm_kernel_launches", "gtsfm_ctx_set_float_screening",
    "gtsfm_store_clear", "gtsfm_image_upload", "gtsfm_image_upload_u8", "gtsfm_match_pair_u8", "gtsfm_match_batch", "gtsfm_match_fetch", "gtsfm_ransac_samples", "gtsfm_ransac_samples_k",
    "gtsfm_verify_batch", "gtsfm_two_view_report_batch", "gtsfm_match_pair", "gtsfm_verify_pair", "gtsfm_debug_fetch",
    "gtsfm_profile_enable", "gtsfm_profile_get", "gtsfm_event_record", "gtsfm_event_elapsed_ms", "gtsfm_verify_stats",
    "gtsfm_host_alloc", "gtsfm_host_free", "gtsfm_match_fetch_async", "gtsfm_sync",
    "gtsfm_retriever_similarity", "gtsfm_retriever_topk", "gtsfm_retriever_pairs_noalloc",
    "gtsfm_verify_hypotheses_used", "gtsfm_result_pack", "gtsfm_result_pack_copy", "gtsfm_stream_wait",
    "gtsfm_images_upload", "gtsfm_two_view_ba_batch",
)

OK, E_CUDA, E_ARG, E_NO_DEVICE, E_KNN_UNPACK, E_UNSUPPORTED = range(6)
V_OK, V_TOO_FEW, V_NO_MODEL = range(3)


class GtsfmB200Error(RuntimeError):
    pass


class VerifyParams(ctypes.Structure):
    _fields_ = [
        ("threshold_px", ctypes.c_double),
        ("n_hypotheses", ctypes.c_int),
        ("lo_iters", ctypes.c_int),
        ("seed", ctypes.c_uint64),
        ("pair_keys", ctypes.c_void_p),
        ("samples", ctypes.c_void_p),
        ("estimate_fundamental", ctypes.c_int),
        ("adaptive_max_hypotheses", ctypes.c_int),
        ("confidence", ctypes.c_double),
    ]


class ReportParams(ctypes.Structure):
    _fields_ = [
        ("dist_threshold_px", ctypes.c_double),
        ("min_num_inliers_est_model", ctypes.c_int),
        ("min_inlier_ratio_est_model", ctypes.c_double),
        ("gt_cal", ctypes.c_void_p),
        ("after_ba", ctypes.c_int),
    ]


class TwoViewBaParams(ctypes.Structure):
    _fields_ = [
        ("min_num_inliers", ctypes.c_int),
        ("max_iterations", ctypes.c_int),
        ("triangulation_reproj_error_thresh", ctypes.c_double),
        ("ba_reproj_error_thresh", ctypes.c_double),
        ("huber_k", ctypes.c_double),
        ("min_triangulation_angle_deg", ctypes.c_double),
    ]


_lib = None
_lib_lock = threading.Lock()


def lib() -> ctypes.CDLL:
    """Load the shared library (built by `make -C gtsfm_b200/csrc` or __graft_entry__.build())."""
    global _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise GtsfmB200Error(
                    f"{LIB_PATH} is missing: build it with `make -C gtsfm_b200/csrc` (nvcc, sm_100a). "
                    "gtsfm_b200 has no CPU fallback."
                )
            L = ctypes.CDLL(LIB_PATH)
            L.gtsfm_last_error.restype = ctypes.c_char_p
            L.gtsfm_last_error.argtypes = [ctypes.c_void_p]
            L.gtsfm_kernel_launches.restype = ctypes.c_int64
            L.gtsfm_kernel_launches.argtypes = [ctypes.c_void_p]
            L.gtsfm_ctx_create.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
            L.gtsfm_ctx_destroy.argtypes = [ctypes.c_void_p]
            L.gtsfm_ctx_destroy.restype = None
            L.gtsfm_ctx_set_workspace.argtypes = [ctypes.c_void_p, ctypes.c_int64]
            L.gtsfm_ctx_set_float_screening.argtypes = [ctypes.c_void_p, ctypes.c_int]
            L.gtsfm_store_clear.argtypes = [ctypes.c_void_p]
            L.gtsfm_host_alloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int64]
            L.gtsfm_host_free.argtypes = [ctypes.c_void_p]
            L.gtsfm_host_free.restype = None
            L.gtsfm_match_fetch_async.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
            L.gtsfm_sync.argtypes = [ctypes.c_void_p]
            L.gtsfm_result_pack.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]
            L.gtsfm_result_pack_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64]
            L.gtsfm_stream_wait.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
            L.gtsfm_verify_hypotheses_used.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
            L.gtsfm_verify_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
            L.gtsfm_image_upload.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_void_p, ctypes.c_void_p]
            L.gtsfm_images_upload.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                              ctypes.c_void_p, ctypes.c_void_p]
            L.gtsfm_image_upload_u8.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                ctypes.c_void_p, ctypes.c_void_p]
            L.gtsfm_match_pair_u8.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int32)]
            L.gtsfm_match_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_void_p,
                                            ctypes.POINTER(ctypes.c_int64)]
            L.gtsfm_match_fetch.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
            L.gtsfm_ransac_samples.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
            L.gtsfm_ransac_samples_k.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                 ctypes.c_void_p]
            L.gtsfm_verify_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.POINTER(VerifyParams)] + [ctypes.c_void_p] * 6
            L.gtsfm_retriever_similarity.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
            L.gtsfm_retriever_topk.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_void_p,
                                               ctypes.c_void_p]
            L.gtsfm_retriever_pairs_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
            L.gtsfm_two_view_report_batch.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                                      ctypes.POINTER(ReportParams)] + [ctypes.c_void_p] * 8
            L.gtsfm_two_view_ba_batch.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(TwoViewBaParams)] + [ctypes.c_void_p] * 6
            L.gtsfm_match_pair.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                           ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int32)]
            L.gtsfm_verify_pair.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                            ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.POINTER(VerifyParams)] + [ctypes.c_void_p] * 6
            L.gtsfm_debug_fetch.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64]
            L.gtsfm_profile_enable.argtypes = [ctypes.c_void_p, ctypes.c_int]
            L.gtsfm_profile_get.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
            L.gtsfm_event_record.argtypes = [ctypes.c_void_p, ctypes.c_int]
            L.gtsfm_event_elapsed_ms.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
            _lib = L
    return _lib


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Context:
    """One CUDA device context of the library (not thread-safe; create one per thread / process)."""

    def __init__(self, device: int = 0):
        self._L = lib()
        h = ctypes.c_void_p()
        rc = self._L.gtsfm_ctx_create(int(device), ctypes.byref(h))
        if rc != OK:
            raise GtsfmB200Error(f"gtsfm_ctx_create failed ({rc}): {self._L.gtsfm_last_error(None).decode()}")
        self._h = h
        self._keep = []   # host arrays whose asynchronous upload may still be in flight
        self._arena = {}  # name -> (pinned pointer, capacity in bytes): result buffers of the reuse_outputs mode
        self._retired = []  # outgrown arena blocks (views handed out earlier stay valid memory until close())
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            for ptr, _ in self._arena.values():
                self._L.gtsfm_host_free(ptr)
            for ptr in self._retired:
                self._L.gtsfm_host_free(ptr)
            self._arena = {}
            self._retired = []
            self._L.gtsfm_ctx_destroy(self._h)
            self._h = None

    def _pinned(self, name: str, shape, dtype) -> np.ndarray:
        """A view of the context's pinned result arena (grown on demand, never shrunk).  Valid until the next call that
        asks for the same name: the reuse_outputs mode trades ownership for full-speed, asynchronous D2H copies."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        nbytes = max(n * dtype.itemsize, 1)
        ptr, cap = self._arena.get(name, (None, 0))
        if cap < nbytes:
            if ptr:
                self._retired.append(ptr)     # arrays handed out earlier may still view it: freed at close(), never before
            cap = int(nbytes * 1.25) + 4096
            p = ctypes.c_void_p()
            self._check(self._L.gtsfm_host_alloc(ctypes.byref(p), cap))
            ptr = p.value
            self._arena[name] = (ptr, cap)
        buf = (ctypes.c_char * nbytes).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)

    def sync(self):
        """Wait for all kernels and asynchronous copies of this context."""
        self._check(self._L.gtsfm_sync(self._h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc == OK:
            return
        msg = self._L.gtsfm_last_error(self._h).decode()
        if rc == E_KNN_UNPACK:
            # the reference's own exception type and text (twoway_matcher.py:137)
            raise ValueError("not enough values to unpack (expected 2, got 1)")
        if rc == E_UNSUPPORTED:
            raise NotImplementedError(msg)
        raise GtsfmB200Error(f"libgtsfm_b200 error {rc}: {msg}")

    @property
    def kernel_launches(self) -> int:
        return int(self._L.gtsfm_kernel_launches(self._h))

    def result_pack(self, n_pairs: int, pair_ids: Optional[np.ndarray] = None) -> int:
        """Build the 128-byte result headers and the payload offset table of the last verify_batch on the device
        (gtsfm_result_pack); returns the number of payload rows (verified correspondences of all pairs)."""
        ids = None if pair_ids is None else np.ascontiguousarray(pair_ids, np.int64)
        rows = ctypes.c_int64(0)
        self._check(self._L.gtsfm_result_pack(self._h, int(n_pairs), _ptr(ids), ctypes.byref(rows)))
        return int(rows.value)

    def result_pack_copy(self, d_dst: int, dst_bytes: int, payload_offset_bytes: int):
        """Write [headers | payload] into a caller-owned DEVICE buffer (address d_dst), on the context's stream."""
        self._check(self._L.gtsfm_result_pack_copy(self._h, ctypes.c_void_p(int(d_dst)), int(dst_bytes), int(payload_offset_bytes)))

    def stream_wait(self, cuda_stream: int, context_waits: bool = False):
        """Order the context's stream against another CUDA stream (a torch stream's .cuda_stream) without a host sync."""
        self._check(self._L.gtsfm_stream_wait(self._h, ctypes.c_void_p(int(cuda_stream)), 1 if context_waits else 0))

    def verify_hypotheses_used(self, n_pairs: int) -> np.ndarray:
        """Hypotheses evaluated per pair by the last verify_batch (differs from n_hypotheses only under the adaptive F budget)."""
        out = np.zeros(int(n_pairs), np.int32)
        self._check(self._L.gtsfm_verify_hypotheses_used(self._h, _ptr(out), int(n_pairs)))
        return out

    def verify_stats(self, reset: bool = False):
        """(candidates scored, hypotheses solved, algorithmic Sampson evaluations, executed evaluations) since the last reset."""
        out = (ctypes.c_int64 * 4)()
        self._check(self._L.gtsfm_verify_stats(self._h, out, 1 if reset else 0))
        return int(out[0]), int(out[1]), int(out[2]), int(out[3])

    def set_float_screening(self, on: bool):
        """Float descriptors with the ratio test off: fp16 tensor-core screening (default) or the exhaustive fp32 kernel."""
        self._check(self._L.gtsfm_ctx_set_float_screening(self._h, 1 if on else 0))

    def set_workspace(self, nbytes: int):
        self._check(self._L.gtsfm_ctx_set_workspace(self._h, int(nbytes)))

    # ---- store
    def clear(self):
        if self._keep:          # uploads nothing has waited for yet may still be reading these host arrays
            self.sync()
        self._check(self._L.gtsfm_store_clear(self._h))
        self._keep.clear()
        self._n_images = 0

    def upload_image(self, image_id: int, desc: Optional[np.ndarray], kps: Optional[np.ndarray] = None, cal=None):
        """desc (n, dim) float32 or None; kps (n, 2) float64 pixel coordinates or None; cal = (f, k1, k2, u0, v0)."""
        n = dim = 0
        if desc is not None and desc.size:
            desc = np.ascontiguousarray(desc, np.float32)
            if desc.ndim != 2:
                raise ValueError("descriptors must be (n, dim)")
            n, dim = desc.shape
        else:
            desc = None
        if kps is not None:
            kps = np.ascontiguousarray(kps, np.float64).reshape(-1, 2)
            if desc is not None and kps.shape[0] != n:
                raise ValueError("keypoints and descriptors disagree on n")
            n = kps.shape[0]
        calv = None if cal is None else np.ascontiguousarray(cal, np.float64)
        self._keep += [desc, kps, calv]
        self._check(self._L.gtsfm_image_upload(self._h, int(image_id), _ptr(desc), int(n), int(dim), _ptr(kps), _ptr(calv)))

    def upload_images(self, first_image_id: int, descs, kps=None, cals=None):
        """A list of float32 descriptor arrays (n_i, dim) [+ (n_i, 2) float64 keypoints, + (f, k1, k2, u0, v0) per image] in one
        library call (gtsfm_images_upload).  The copies are enqueued on the context's copy stream; the arrays are kept alive
        until clear()."""
        descs = [np.ascontiguousarray(d, np.float32) for d in descs]
        m = len(descs)
        if m == 0:
            return
        dim = int(descs[0].shape[1])
        if any(d.ndim != 2 or d.shape[1] != dim for d in descs):
            raise ValueError("descriptors must be (n, dim) arrays of one dimension")
        n = np.array([d.shape[0] for d in descs], np.int32)
        dptr = (ctypes.c_void_p * m)(*[d.ctypes.data for d in descs])
        kptr = None
        if kps is not None:
            kps = [np.ascontiguousarray(k, np.float64).reshape(-1, 2) for k in kps]
            if any(k.shape[0] != d.shape[0] for k, d in zip(kps, descs)):
                raise ValueError("keypoints and descriptors disagree on n")
            kptr = (ctypes.c_void_p * m)(*[k.ctypes.data for k in kps])
        calv = None if cals is None else np.ascontiguousarray(cals, np.float64).reshape(m, 5)
        self._keep += [descs, kps, calv, n]
        self._check(self._L.gtsfm_images_upload(self._h, int(first_image_id), m, dptr, _ptr(n), dim, kptr, _ptr(calv)))

    def upload_image_u8(self, image_id: int, desc: np.ndarray, kps: Optional[np.ndarray] = None, cal=None):
        """Binary descriptors (n, nbytes) uint8 for the HAMMING distance; kps / cal as in upload_image."""
        if desc is None or np.size(desc) == 0:      # an image without features: its pairs yield np.array([]) (twoway_matcher.py:71-72)
            desc, n, nbytes = None, 0, 0
        else:
            desc = np.ascontiguousarray(desc, np.uint8)
            if desc.ndim != 2:
                raise ValueError("descriptors must be (n, nbytes)")
            n, nbytes = desc.shape
        if kps is not None:
            kps = np.ascontiguousarray(kps, np.float64).reshape(-1, 2)
            if kps.shape[0] != n:
                raise ValueError("keypoints and descriptors disagree on n")
            if n == 0:
                kps = None
        calv = None if cal is None else np.ascontiguousarray(cal, np.float64)
        self._keep += [desc, kps, calv]
        self._check(self._L.gtsfm_image_upload_u8(self._h, int(image_id), _ptr(desc), int(n), int(nbytes), _ptr(kps), _ptr(calv)))

    # ---- matcher
    def match_batch(self, pairs: np.ndarray, ratio: Optional[float], fetch=True, reuse_outputs: bool = False):
        """Returns (matches (total, 2) uint32, offsets (P + 1) int64); fetch=False leaves the matches on the device.
        fetch="async" (implies reuse_outputs) only enqueues the copy, which then overlaps verify_batch: the returned
        array is a view of the context's pinned arena and must not be read before sync()."""
        pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        P = pairs.shape[0]
        counts = np.zeros(P, np.int32)
        total = ctypes.c_int64(0)
        r = float("nan") if ratio is None else float(ratio)
        self._check(self._L.gtsfm_match_batch(self._h, _ptr(pairs), P, r, _ptr(counts), ctypes.byref(total)))
        # uploads of images this batch does not name may still be in flight: their host arrays stay referenced until clear()
        if fetch == "async":
            out = self._pinned("matches", (total.value, 2), np.uint32)
            self._check(self._L.gtsfm_match_fetch_async(self._h, _ptr(out)))
        else:
            shape = (total.value if fetch else 0, 2)
            out = self._pinned("matches", shape, np.uint32) if (reuse_outputs and fetch) else np.empty(shape, np.uint32)
            if fetch:
                self._check(self._L.gtsfm_match_fetch(self._h, _ptr(out)))
        offsets = np.zeros(P + 1, np.int64)
        np.cumsum(counts, out=offsets[1:])
        return out, offsets

    def match_pair(self, d1: np.ndarray, d2: np.ndarray, ratio: Optional[float]) -> np.ndarray:
        d1 = np.ascontiguousarray(d1, np.float32)
        d2 = np.ascontiguousarray(d2, np.float32)
        if d1.size == 0 or d2.size == 0:
            return np.empty((0, 2), np.uint32)
        n1, dim = d1.shape
        n2 = d2.shape[0]
        out = np.empty((min(n1, n2), 2), np.uint32)
        cnt = ctypes.c_int32(0)
        r = float("nan") if ratio is None else float(ratio)
        self._check(self._L.gtsfm_match_pair(self._h, _ptr(d1), n1, _ptr(d2), n2, dim, r, _ptr(out), ctypes.byref(cnt)))
        return out[: cnt.value].copy()

    def match_pair_u8(self, d1: np.ndarray, d2: np.ndarray, ratio: Optional[float]) -> np.ndarray:
        """TwoWayMatcher(HAMMING).match for one pair of (n, nbytes) uint8 arrays."""
        d1 = np.ascontiguousarray(d1, np.uint8)
        d2 = np.ascontiguousarray(d2, np.uint8)
        if d1.size == 0 or d2.size == 0:
            return np.empty((0, 2), np.uint32)
        n1, nbytes = d1.shape
        n2 = d2.shape[0]
        out = np.empty((min(n1, n2), 2), np.uint32)
        cnt = ctypes.c_int32(0)
        r = float("nan") if ratio is None else float(ratio)
        self._check(self._L.gtsfm_match_pair_u8(self._h, _ptr(d1), n1, _ptr(d2), n2, nbytes, r, _ptr(out), ctypes.byref(cnt)))
        return out[: cnt.value].copy()

    PROFILE_SLOTS = ("prepare", "screen", "resolve_rows", "resolve_cols", "exact_scan", "match_finalize", "gather",
                     "solve", "score", "verify_finalize")

    def profile_enable(self, on: bool = True):
        self._check(self._L.gtsfm_profile_enable(self._h, int(bool(on))))

    def profile_get(self) -> dict:
        """{kernel: (milliseconds, launches)} accumulated since profile_enable(True); CUDA-event timed on the device."""
        n = len(self.PROFILE_SLOTS)
        ms = np.zeros(n, np.float64)
        calls = np.zeros(n, np.int64)
        self._check(self._L.gtsfm_profile_get(self._h, _ptr(ms), _ptr(calls), n))
        return {k: (float(ms[i]), int(calls[i])) for i, k in enumerate(self.PROFILE_SLOTS)}

    def event_record(self, which: int):
        self._check(self._L.gtsfm_event_record(self._h, int(which)))

    def event_elapsed_ms(self, a: int, b: int) -> float:
        ms = ctypes.c_double(0)
        self._check(self._L.gtsfm_event_elapsed_ms(self._h, int(a), int(b), ctypes.byref(ms)))
        return ms.value

    def debug_fetch(self, what: int, count: int, dtype) -> np.ndarray:
        out = np.zeros(count, dtype)
        self._check(self._L.gtsfm_debug_fetch(self._h, int(what), _ptr(out), out.nbytes))
        return out

    # ---- verifier
    def verify_batch(self, pairs: np.ndarray, params: "VerifyParams", match_idx: Optional[np.ndarray] = None,
                     match_offsets: Optional[np.ndarray] = None, total: Optional[int] = None, reuse_outputs: bool = False,
                     fetch: bool = True):
        """Verify the last match batch (match_idx None) or caller-supplied matches.  Returns a dict of arrays; with
        reuse_outputs they are views of the context's pinned arena, valid until the next verify_batch.  fetch=False leaves
        every result on the device (for result_pack / two_view_report_batch) and returns None."""
        pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        P = pairs.shape[0]
        if match_idx is not None:
            match_idx = np.ascontiguousarray(match_idx, np.uint32).reshape(-1, 2)
            match_offsets = np.ascontiguousarray(match_offsets, np.int64)
            total = int(match_offsets[-1])
        if not fetch:
            self._check(self._L.gtsfm_verify_batch(self._h, _ptr(pairs), P, _ptr(match_idx), _ptr(match_offsets), ctypes.byref(params),
                                                   None, None, None, None, None, None))
            return None
        if reuse_outputs:
            R = self._pinned("R", (P, 9), np.float64); t = self._pinned("t", (P, 3), np.float64); E = self._pinned("E", (P, 9), np.float64)
            ninl = self._pinned("ninl", (P,), np.int32); status = self._pinned("status", (P,), np.int32)
            mask = self._pinned("mask", (max(int(total or 0), 1),), np.uint8)
        else:
            R = np.zeros((P, 9)); t = np.zeros((P, 3)); E = np.zeros((P, 9))
            ninl = np.zeros(P, np.int32); status = np.zeros(P, np.int32)
            mask = np.zeros(max(int(total or 0), 1), np.uint8)
        self._check(self._L.gtsfm_verify_batch(self._h, _ptr(pairs), P, _ptr(match_idx), _ptr(match_offsets), ctypes.byref(params),
                                               _ptr(R), _ptr(t), _ptr(ninl), _ptr(status), _ptr(mask), _ptr(E)))
        return dict(R=R.reshape(P, 3, 3), t=t, n_inliers=ninl, status=status, mask=mask[: int(total or 0)], E=E.reshape(P, 3, 3))

    def retriever_similarity(self, global_desc: np.ndarray, blocksize: int = 50, fetch: bool = True) -> Optional[np.ndarray]:
        """sim = G G^T of (n_images, dim) float32 global descriptors (gtsfm_retriever_similarity); the matrix also stays on
        the device for retriever_topk(None, ...)."""
        g = np.ascontiguousarray(global_desc, np.float32)
        if g.ndim != 2:
            raise ValueError("global descriptors must be (n_images, dim)")
        n, dim = g.shape
        sim = np.empty((n, n), np.float32) if fetch else None
        self._check(self._L.gtsfm_retriever_similarity(self._h, _ptr(g), n, dim, int(blocksize), _ptr(sim)))
        return sim

    def retriever_topk(self, sim: Optional[np.ndarray], n_images: int, num_select: int, min_score: Optional[float]):
        """(idx (n, k) int32 with -1 for unused slots, score (n, k) float32), k = min(num_select, n) (gtsfm_retriever_topk)."""
        n = int(n_images)
        k = min(int(num_select), n)
        if sim is not None:
            sim = np.ascontiguousarray(sim, np.float32)
            if sim.shape != (n, n):
                raise ValueError("similarity matrix must be (n_images, n_images)")
        idx = np.empty((n, k), np.int32)
        score = np.empty((n, k), np.float32)
        ms = float("nan") if min_score is None else float(min_score)
        self._check(self._L.gtsfm_retriever_topk(self._h, _ptr(sim), n, k, ms, _ptr(idx), _ptr(score)))
        return idx, score

    def two_view_report_batch(self, n_pairs: int, total: int, dist_threshold_px: float, min_num_inliers_est_model: int,
                              min_inlier_ratio_est_model: float, gt_i2Ri1: Optional[np.ndarray] = None,
                              gt_i2ti1: Optional[np.ndarray] = None, reuse_outputs: bool = False, gt_cal: Optional[np.ndarray] = None,
                              after_ba: bool = False):
        """Post-verification report of the pairs of the last verify_batch (gtsfm_two_view_report_batch).  With reuse_outputs the
        per-correspondence arrays are views of the context's pinned arena (valid until the next report call).  after_ba: the
        report of the results two_view_ba_batch left on the device (the post-BA report)."""
        P = int(n_pairs)
        if gt_cal is not None:
            gt_cal = np.ascontiguousarray(gt_cal, np.float64).reshape(P, 10)
        prm = ReportParams(float(dist_threshold_px), int(min_num_inliers_est_model), float(min_inlier_ratio_est_model),
                           None if gt_cal is None else gt_cal.ctypes.data, 1 if after_ba else 0)
        if gt_i2Ri1 is not None:
            gt_i2Ri1 = np.ascontiguousarray(gt_i2Ri1, np.float64).reshape(P, 9)
            gt_i2ti1 = np.ascontiguousarray(gt_i2ti1, np.float64).reshape(P, 3)
        nt = max(int(total), 1)                      # the kernel writes every row (NaN / 0 where nothing was verified)
        dist = self._pinned("report_dist", (nt,), np.float64) if reuse_outputs else np.empty(nt, np.float64)
        gmask = self._pinned("report_mask", (nt,), np.uint8) if reuse_outputs else np.empty(nt, np.uint8)
        if int(total) == 0:
            dist[:] = np.nan; gmask[:] = 0
        n_gt = np.zeros(P, np.int32); keep = np.zeros(P, np.int32)
        in_avg = np.zeros(P); out_avg = np.zeros(P); r_err = np.zeros(P); u_err = np.zeros(P)
        self._check(self._L.gtsfm_two_view_report_batch(self._h, P, _ptr(gt_i2Ri1), _ptr(gt_i2ti1), ctypes.byref(prm), _ptr(dist),
                                                        _ptr(gmask), _ptr(n_gt), _ptr(in_avg), _ptr(out_avg), _ptr(r_err),
                                                        _ptr(u_err), _ptr(keep)))
        return dict(dist_sq=dist[: int(total)], gt_mask=gmask[: int(total)], num_inliers_gt=n_gt, inlier_avg_err=in_avg,
                    outlier_avg_err=out_avg, R_error_deg=r_err, U_error_deg=u_err, keep=keep)

    def two_view_ba_batch(self, n_pairs: int, total: int, min_num_inliers: int, max_iterations: int = 100,
                          triangulation_reproj_error_thresh: float = 100.0, ba_reproj_error_thresh: float = 0.5, huber_k: float = 1.345,
                          min_triangulation_angle_deg: float = 0.0):
        """Triangulation + two-view bundle adjustment of the pairs of the last verify_batch (gtsfm_two_view_ba_batch).  Returns
        R (P, 3, 3), t (P, 3) unit, valid_mask (total,) uint8 over the match rows, num_valid, status (0 ran / 1 skipped / 2
        nothing triangulated), iterations."""
        P = int(n_pairs)
        prm = TwoViewBaParams(int(min_num_inliers), int(max_iterations), float(triangulation_reproj_error_thresh),
                              float(ba_reproj_error_thresh), float(huber_k), float(min_triangulation_angle_deg))
        R = np.zeros((P, 3, 3)); t = np.zeros((P, 3))
        mask = np.zeros(max(int(total), 1), np.uint8)
        n = np.zeros(P, np.int32); status = np.zeros(P, np.int32); iters = np.zeros(P, np.int32)
        self._check(self._L.gtsfm_two_view_ba_batch(self._h, P, ctypes.byref(prm), _ptr(R), _ptr(t), _ptr(mask), _ptr(n), _ptr(status),
                                                    _ptr(iters)))
        return dict(R=R, t=t, valid_mask=mask[: int(total)], num_valid=n, status=status, iterations=iters)

    def verify_pair(self, kps1, kps2, match_idx, cal1, cal2, params: "VerifyParams"):
        kps1 = np.ascontiguousarray(kps1, np.float64).reshape(-1, 2)
        kps2 = np.ascontiguousarray(kps2, np.float64).reshape(-1, 2)
        mi = np.ascontiguousarray(match_idx, np.uint32).reshape(-1, 2)
        c1 = np.ascontiguousarray(cal1, np.float64); c2 = np.ascontiguousarray(cal2, np.float64)
        M = mi.shape[0]
        R = np.zeros(9); t = np.zeros(3); E = np.zeros(9)
        ninl = ctypes.c_int32(0); status = ctypes.c_int32(0)
        mask = np.zeros(max(M, 1), np.uint8)
        self._check(self._L.gtsfm_verify_pair(self._h, _ptr(kps1), kps1.shape[0], _ptr(kps2), kps2.shape[0], _ptr(mi), M, _ptr(c1),
                                              _ptr(c2), ctypes.byref(params), _ptr(R), _ptr(t), ctypes.addressof(ninl),
                                              ctypes.addressof(status), _ptr(mask), _ptr(E)))
        return dict(R=R.reshape(3, 3), t=t, n_inliers=ninl.value, status=status.value, mask=mask[:M], E=E.reshape(3, 3))


def ransac_samples(seed: int, pair_key: int, n_matches: int, n_hyp: int, sample_size: int = 5) -> np.ndarray:
    """The (n_hyp, sample_size) sample-index set the device draws for (seed, pair_key): feed it to the CPU oracle.
    sample_size 5 = essential-matrix mode, 7 = fundamental-matrix mode."""
    out = np.zeros((n_hyp, sample_size), np.int32)
    rc = lib().gtsfm_ransac_samples_k(int(seed), int(pair_key), int(n_matches), int(n_hyp), int(sample_size), _ptr(out))
    if rc != OK:
        raise GtsfmB200Error(f"gtsfm_ransac_samples_k failed ({rc})")
    return out


DEFAULT_PLUGIN_WORKSPACE_GIB = 4.0
_default_tls = threading.local()


def default_context(device: int = 0) -> Context:
    """Lazily created context of the calling THREAD of this process for `device`.  Plugin objects never hold one, so they
    stay picklable (det_desc_correspondence_generator.py:67 scatters them); a context is single-threaded, so every thread
    of a Dask worker (--threads_per_worker > 1, gtsfm_runner_base.py:58-69) gets its own store, stream and work buffers:
    match() / verify() are re-entrant like the reference's, which builds a fresh cv2 matcher per call
    (twoway_matcher.py:133).  A context created before a fork is never reused in the child (key includes the pid)."""
    table = getattr(_default_tls, "ctx", None)
    if table is None:
        table = _default_tls.ctx = {}
    key = (os.getpid(), int(device))
    if key not in table:
        ctx = Context(device)
        # The library sizes its per-chunk work buffers by this budget (they grow on demand up to it and are kept).  Contexts of
        # this kind exist once per worker THREAD, so the plugin default is 4 GiB instead of the library's 12 GiB; a smaller
        # budget only means more, smaller chunks.  GTSFM_B200_WORKSPACE_GIB overrides it.
        ctx.set_workspace(int(float(os.environ.get("GTSFM_B200_WORKSPACE_GIB", DEFAULT_PLUGIN_WORKSPACE_GIB)) * (1 << 30)))
        table[key] = ctx
    return table[key]
