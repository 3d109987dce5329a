"""b200attn -- host side of the B200 paged-attention decode backend.

`ops` are the two device ops (KV write, paged decode) over libb200attn.so;
`BlockManager` is the C++ page allocator.  The CUDA extension is mandatory.
"""
from ._lib import B2Error, KV_BF16, KV_FP8_E4M3, LIB_PATH, load  # noqa: F401
from .block_manager import BlockManager  # noqa: F401
from .ops import (append_is_fused, decode_auto_splits, decode_workspace_bytes, paged_decode,  # noqa: F401
                  paged_decode_append, plan_num_splits, reshape_and_cache)
from .step import DecodeStep  # noqa: F401,E402
from .scheduler import Scheduler, StepPlan  # noqa: F401,E402
