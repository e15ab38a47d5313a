"""ctypes binding of lib/libevr.so -- the C ABI declared in include/evr.h."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EVR_LIB_PATH") or os.path.join(_HERE, "lib", "libevr.so")  # override: A/B builds


class EvrError(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python __graft_entry__.py` (nvcc, sm_100a). "
        "This package has no CPU fallback.")

lib = ctypes.CDLL(LIB_PATH)

c_void_p, c_int, c_int64, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t


class VitCfg(ctypes.Structure):
    _fields_ = [("image_size", c_int), ("patch", c_int), ("width", c_int), ("layers", c_int),
                ("heads", c_int), ("out_dim", c_int)]


class TextCfg(ctypes.Structure):
    _fields_ = [("context_length", c_int), ("vocab_size", c_int), ("width", c_int), ("layers", c_int),
                ("heads", c_int), ("out_dim", c_int)]


# name -> (restype, argtypes); every symbol include/evr.h declares
SIGNATURES = {
    "evr_last_error": (ctypes.c_char_p, []),
    "evr_version": (c_int, []),
    "evr_launch_count": (c_int64, []),
    "evr_profile_enable": (c_int, [c_int]),
    "evr_profile_read": (c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_int64), c_int]),
    "evr_profile_timeline": (c_int, [ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(c_int), c_int,
                                     ctypes.POINTER(c_int)]),
    "evr_store_prepare": (c_int, [c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p]),
    "evr_topk_workspace_bytes": (c_int, [c_int64, c_int, c_int, c_int, ctypes.POINTER(c_size_t)]),
    "evr_topk_search": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                c_void_p, c_size_t, c_void_p]),
    "evr_topk_search_static": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                       c_void_p]),
    "evr_topk_merge": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "evr_topk_large_workspace_bytes": (c_int, [c_int64, c_int, c_int64, ctypes.POINTER(c_size_t)]),
    "evr_topk_search_large": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_int, c_int64, c_void_p, c_void_p,
                                      c_void_p, c_size_t, c_void_p]),
    "evr_topk_merge_large": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_size_t,
                                     c_void_p]),
    "evr_rescore": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "evr_vit_create": (c_int, [ctypes.POINTER(VitCfg), ctypes.POINTER(c_void_p), c_int, c_int,
                               ctypes.POINTER(c_void_p)]),
    "evr_vit_destroy": (c_int, [c_void_p]),
    "evr_vit_encode": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "evr_resize_crop_u8": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                   c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "evr_scene_deltas": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "evr_text_create": (c_int, [ctypes.POINTER(TextCfg), ctypes.POINTER(c_void_p), c_int, c_int,
                                ctypes.POINTER(c_void_p)]),
    "evr_text_destroy": (c_int, [c_void_p]),
    "evr_text_encode": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p]),
    "evr_jpeg_create": (c_int, [c_int, ctypes.POINTER(c_void_p)]),
    "evr_jpeg_destroy": (c_int, [c_void_p]),
    "evr_jpeg_backend": (c_int, [c_void_p]),
    "evr_jpeg_info": (c_int, [c_void_p, c_void_p, c_size_t, ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                              ctypes.POINTER(c_int)]),
    "evr_jpeg_decode_batch": (c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_size_t), c_int, c_int, c_int,
                                      c_void_p, c_void_p]),
    "evr_layernorm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "evr_attention": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "evr_gemm_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
}
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


def check(status, what):
    if status != 0:
        raise EvrError(f"{what} failed ({status}): {lib.evr_last_error().decode(errors='replace')}")


def launch_count():
    return int(lib.evr_launch_count())


PROFILE_FAMILIES = ("gemm_tc", "attention", "layernorm", "vit_misc", "topk_scan", "topk_final", "store")


def profile_enable(on):
    """True / 1: CUDA events around every launch (all families).  2: no events, the GEMM / LayerNorm / attention
    kernels stamp %globaltimer themselves (timed in place, PDL overlap intact).  False / 0: off."""
    check(lib.evr_profile_enable(2 if on == 2 else (1 if on else 0)), "evr_profile_enable")


def profile_read():
    """-> {family: (total_ms, launches)} since profile_enable(True)."""
    n = len(PROFILE_FAMILIES)
    ms = (ctypes.c_double * n)()
    cnt = (c_int64 * n)()
    check(lib.evr_profile_read(ms, cnt, n), "evr_profile_read")
    return {f: (ms[i], int(cnt[i])) for i, f in enumerate(PROFILE_FAMILIES)}


def profile_timeline(capacity=16384):
    """After profile_enable(2) and some launches: [(family, start_ns, end_ns)] in launch order (%globaltimer)."""
    se = (ctypes.c_uint64 * (2 * capacity))()
    fam = (c_int * capacity)()
    n = c_int(0)
    check(lib.evr_profile_timeline(se, fam, capacity, ctypes.byref(n)), "evr_profile_timeline")
    return [(PROFILE_FAMILIES[fam[i]], int(se[2 * i]), int(se[2 * i + 1])) for i in range(n.value)
            if se[2 * i + 1] > se[2 * i]]


def current_stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    """Device (or host) pointer of a torch tensor as c_void_p; None -> NULL."""
    return c_void_p(0 if t is None else t.data_ptr())


class HandleGuard:
    """Serialises use of one native handle's device workspace across host threads AND CUDA streams.

    The reference runs under a threaded Flask server and holds no locks; a handle here owns one set of
    activation / candidate buffers, so two calls must not overlap on the device even when they come from different
    streams: the guard takes a host lock for the enqueue and makes the caller's stream wait for the previous call's
    completion event before the kernels that reuse the workspace are launched."""

    def __init__(self):
        import threading
        self._lock = threading.Lock()
        self._done = None

    def __enter__(self):
        import torch
        self._lock.acquire()
        if self._done is not None:
            torch.cuda.current_stream().wait_event(self._done)
        return self

    def __exit__(self, *exc):
        import torch
        try:
            if self._done is None:
                self._done = torch.cuda.Event()
            self._done.record(torch.cuda.current_stream())
        finally:
            self._lock.release()
        return False
