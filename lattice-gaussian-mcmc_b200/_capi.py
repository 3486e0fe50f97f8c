"""ctypes binding of liblgs_b200.so (include/lgs_b200.h).  Fails loudly when the library is missing: there is no
Python or CPU stand-in for the sampling kernels."""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# LGS_B200_LIB: development override used by tools/ to A/B kernel variants; the default is the in-tree build
LIB_PATH = os.environ.get("LGS_B200_LIB") or os.path.join(HERE, "liblgs_b200.so")

LGS_OK = 0
LGS_E_INVALID, LGS_E_CUDA, LGS_E_NOMEM, LGS_E_STATE, LGS_E_FALLBACK = -1, -2, -3, -4, -5
LGS_E_RANGE, LGS_E_PRECISION = -6, -7
SAMPLEZ = {"table": 0, "fast": 1}
WEIGHT = {"reference": 0, "exact": 1}
PATH = {"auto": 0, "ordered": 1, "blocked": 2}

c_f64p = ctypes.POINTER(ctypes.c_double)
c_i64p = ctypes.POINTER(ctypes.c_int64)
c_i32p = ctypes.POINTER(ctypes.c_int32)
c_u8p = ctypes.POINTER(ctypes.c_uint8)
vp = ctypes.c_void_p


class LgsDesc(ctypes.Structure):
    _fields_ = [
        ("n", ctypes.c_int32),
        ("R", c_f64p),
        ("B", c_f64p),
        ("t", c_f64p),
        ("center", c_f64p),
        ("sigma", ctypes.c_double),
        ("precision", ctypes.c_int32),
        ("max_range", ctypes.c_int32),
        ("device", ctypes.c_int32),
        ("path", ctypes.c_int32),
    ]


class LgsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"liblgs_b200 error {code}: {msg}")
        self.code = code


_lib = None

# name -> (restype, argtypes); mirrors include/lgs_b200.h one to one
PROTOTYPES = {
    "lgs_abi_version": (ctypes.c_int, []),
    "lgs_create": (ctypes.c_int, [ctypes.POINTER(LgsDesc), ctypes.POINTER(vp)]),
    "lgs_destroy": (None, [vp]),
    "lgs_last_error": (ctypes.c_char_p, [vp]),
    "lgs_uniforms_per_sample": (ctypes.c_int, [vp]),
    "lgs_preferred_batch": (ctypes.c_int64, [vp]),
    "lgs_klein_sample": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, vp, vp, vp, vp]),
    "lgs_klein_replay": (ctypes.c_int, [vp, ctypes.c_int64, vp, vp, vp, vp, vp, vp]),
    "lgs_replay_path": (ctypes.c_int, [vp]),
    "lgs_host_transport": (ctypes.c_int, [vp]),
    "lgs_host_widen_i16": (ctypes.c_int, [vp, vp, ctypes.c_int64, ctypes.c_int32]),
    "lgs_klein_sample_host": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, vp, vp, vp]),
    "lgs_klein_replay_host": (ctypes.c_int, [vp, ctypes.c_int64, vp, vp, vp, vp, vp]),
    "lgs_klein_logdensity": (ctypes.c_int, [vp, ctypes.c_int64, vp, vp, vp]),
    "lgs_klein_logdensity_host": (ctypes.c_int, [vp, ctypes.c_int64, vp, vp]),
    "lgs_imhk_reset": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint64]),
    "lgs_imhk_run": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int, vp, vp, vp]),
    "lgs_imhk_replay": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, vp, ctypes.c_int64, vp, vp, vp, vp, vp, vp]),
    "lgs_imhk_get_state": (ctypes.c_int, [vp, vp, vp, vp, vp, vp]),
    "lgs_imhk_run_host": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int, vp, vp]),
    "lgs_imhk_replay_host": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, vp, ctypes.c_int64, vp, vp, vp, vp, vp]),
    "lgs_imhk_get_state_host": (ctypes.c_int, [vp, vp, vp, vp, vp]),
    "lgs_imhk_set_state": (ctypes.c_int, [vp, ctypes.c_int, vp, vp, vp, ctypes.c_int64, vp]),
    "lgs_imhk_set_state_host": (ctypes.c_int, [vp, ctypes.c_int, vp, vp, vp, ctypes.c_int64]),
    "lgs_imhk_steps_done": (ctypes.c_int64, [vp]),
    "lgs_imhk_counts_host": (ctypes.c_int, [vp, c_i64p, c_i64p]),
    "lgs_chain_diagnostics": (ctypes.c_int, [vp, vp, ctypes.c_int64, ctypes.c_int64, c_i32p, ctypes.c_int32, ctypes.c_double,
                                             ctypes.c_int32, vp, vp, ctypes.c_int32, vp]),
    "lgs_chain_diagnostics_static": (ctypes.c_int, [vp, vp, ctypes.c_int64, ctypes.c_int64, c_i32p, ctypes.c_int32,
                                                    ctypes.c_double, ctypes.c_int32, vp, vp, ctypes.c_int32, vp]),
    "lgs_gelman_rubin_sums": (ctypes.c_int, [vp, vp, ctypes.c_int64, ctypes.c_int32, vp, vp]),
    "lgs_lll_reduce": (ctypes.c_int, [vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_double, ctypes.c_double, ctypes.c_int32, vp, vp]),
    "lgs_bkz_reduce": (ctypes.c_int, [vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_double, ctypes.c_int32,
                                      ctypes.c_int32, ctypes.c_int32, vp, vp]),
    "lgs_gso_norms": (ctypes.c_int, [vp, ctypes.c_int32, ctypes.c_int32, vp]),
    "lgs_profile_enable": (ctypes.c_int, [vp, ctypes.c_int]),
    "lgs_profile_read": (ctypes.c_int, [vp, c_f64p, c_i64p]),
    "lgs_histogram": (ctypes.c_int, [vp, vp, ctypes.c_int64, c_i32p, ctypes.c_int32, ctypes.c_double, ctypes.c_double, ctypes.c_int32, vp, vp]),
    "lgs_attach_histogram": (ctypes.c_int, [vp, c_i32p, ctypes.c_int32, ctypes.c_double, ctypes.c_double, ctypes.c_int32, vp]),
}


def lib():
    """Load liblgs_b200.so; raise (never fall back) if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build the CUDA extension first "
                "(python lattice-gaussian-mcmc_b200/build.py, or __graft_entry__.build()). "
                "There is no CPU fallback for the sampling kernels.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(handle, rc):
    if rc != LGS_OK:
        msg = lib().lgs_last_error(handle)
        raise LgsError(rc, msg.decode() if msg else "")
