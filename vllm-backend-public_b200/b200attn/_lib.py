"""ctypes binding of libb200attn.so (C ABI: include/b200attn.h).

There is deliberately no fallback: if the shared library is missing the import of any op
fails with a clear message (run `python -c "import __graft_entry__ as g; g.build()"` or
`make -C vllm-backend-public_b200/csrc`).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG_ROOT = Path(__file__).resolve().parent.parent          # vllm-backend-public_b200/
LIB_PATH = Path(os.environ.get("B200ATTN_LIB", PKG_ROOT / "lib" / "libb200attn.so"))

B2_OK, B2_EINVAL, B2_EUNSUPPORTED, B2_ECUDA, B2_ENOMEM, B2_ENOTFOUND, B2_EEXISTS = 0, -1, -2, -3, -4, -5, -6
KV_BF16, KV_FP8_E4M3 = 0, 1
ABI_VERSION = 1

_ERRNAMES = {-1: "B2_EINVAL", -2: "B2_EUNSUPPORTED", -3: "B2_ECUDA", -4: "B2_ENOMEM",
             -5: "B2_ENOTFOUND", -6: "B2_EEXISTS"}


class B2Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{_ERRNAMES.get(code, code)}: {msg}")
        self.code = code


vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

# name -> (restype, argtypes): one entry per declaration in include/b200attn.h
SIGNATURES = {
    "b2_version": (i32, []),
    "b2_last_error": (C.c_char_p, []),
    "b2_reshape_and_cache": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i64, i64, i32, vp, vp, vp]),
    "b2_reshape_and_cache_decode": (i32, [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i64, i32,
                                          vp, vp, vp]),
    "b2_paged_decode_workspace_bytes": (sz, [i32, i32, i32, i32, i32, i32, i32]),
    "b2_paged_decode_auto_splits": (i32, [i32, i32, i32, i32, i32, i32, i32]),
    "b2_paged_decode_plan_splits": (i32, [vp, i32, i32, i32, i32, i32, i32]),
    "b2_paged_decode": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, i64, i64,
                              f32, i32, vp, vp, vp, sz, i32, vp]),
    "b2_paged_decode_append_fused": (i32, [i32, i32, i32]),
    "b2_paged_decode_append": (i32, [vp, vp, vp, vp, vp, i64, i64, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i64,
                                     i64, i64, i64, i64, f32, i32, vp, vp, vp, sz, i32, vp]),
    "b2_paged_decode_static": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, i64, i64,
                                     f32, i32, vp, vp, i32, vp]),
    "b2_decode_plan_create": (vp, [vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, f32, i32, vp, vp, vp, sz, vp]),
    "b2_decode_plan_destroy": (None, [vp]),
    "b2_decode_plan_run": (i32, [vp, vp, vp, vp, vp, i32, i64, i64, i32, vp]),
    "b2_decode_plan_run_append": (i32, [vp, vp, vp, vp, vp, vp, i64, i64, vp, vp, i32, i64, i64, i32, vp]),
    "b2_paged_decode_simt": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, i64,
                                   f32, i32, vp, vp, vp]),
    "b2_prefill_attention": (i32, [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i64, i64, i64,
                                   i64, i64, f32, i32, vp, vp, vp]),
    "b2_dequant_kv_blocks": (i32, [vp, vp, vp, vp, vp, i64, i32, i32, i32, i64, i64, vp]),
    "b2_rms_norm": (i32, [vp, vp, vp, vp, i32, i32, i64, i64, i64, f32, vp]),
    "b2_silu_and_mul": (i32, [vp, vp, i32, i32, i64, i64, vp]),
    "b2_rope_and_cache": (i32, [vp, vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, vp, vp, vp]),
    "b2_rope_and_cache_ex": (i32, [vp, vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, vp, vp, i32, vp]),
    "b2_skinny_gemm": (i32, [vp, vp, vp, i32, i32, i32, i64, i64, vp]),
    "b2_sample": (i32, [vp, vp, i32, i32, i64, vp, vp, vp]),
    "b2_sample_filtered": (i32, [vp, vp, i32, i32, i64, vp, vp, vp, vp, vp]),
    "b2_bm_create": (vp, [i64, i32]),
    "b2_bm_destroy": (None, [vp]),
    "b2_bm_num_free": (i64, [vp]),
    "b2_bm_num_seqs": (i64, [vp]),
    "b2_bm_seq_len": (i64, [vp, i64]),
    "b2_bm_can_append": (i32, [vp, i64, i64]),
    "b2_bm_alloc_seq": (i32, [vp, i64, i64]),
    "b2_bm_append": (i32, [vp, i64, i64, vp]),
    "b2_bm_free_seq": (i32, [vp, i64]),
    "b2_bm_fill_tables": (i32, [vp, vp, i32, vp, i32, vp]),
    "b2_bm_plan_decode": (i32, [vp, vp, i32, vp, i32, vp, vp]),
    "b2_sched_create": (vp, [i64, i32, i32, i32, i32]),
    "b2_sched_destroy": (None, [vp]),
    "b2_sched_add_request": (i32, [vp, i64, i32, i32]),
    "b2_sched_finish": (i32, [vp, i64]),
    "b2_sched_step": (i32, [vp, vp]),
    "b2_sched_num_waiting": (i64, [vp]),
    "b2_sched_num_running": (i64, [vp]),
    "b2_sched_num_free_blocks": (i64, [vp]),
    "b2_sched_generated": (i64, [vp, i64]),
    "b2_comm_alloc": (i32, [sz, vp, vp]),
    "b2_comm_open": (i32, [vp, vp]),
    "b2_comm_close": (i32, [vp]),
    "b2_comm_free": (i32, [vp]),
    "b2_comm_region_bytes": (sz, [i32, i32, i32, i32, i32]),
    "b2_comm_create": (vp, [i32, i32, vp, sz, i32, i32, i32, i32]),
    "b2_comm_destroy": (None, [vp]),
    "b2_comm_input": (vp, [vp, i32]),
    "b2_comm_gathered": (vp, [vp]),
    "b2_comm_debug_stamps": (i32, [vp, vp]),
    "b2_comm_allreduce_rms_norm": (i32, [vp, i32, vp, vp, vp, i32, i64, i64, f32, vp]),
    "b2_comm_push_allreduce_rms_norm": (i32, [vp, vp, i64, vp, vp, vp, i32, i64, i64, f32, vp]),
    "b2_comm_allgather_cols": (i32, [vp, vp, i64, i32, i32, vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load the library once; raise loudly when it is absent or its ABI differs."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} not found: the sm_100a extension is not built. b200attn has no CPU or "
            f"PyTorch fallback; build it with `make -C {PKG_ROOT / 'csrc'}`.")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    v = lib.b2_version()
    if v != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {v}, binding expects {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != B2_OK:
        raise B2Error(rc, load().b2_last_error().decode("utf-8", "replace"))
