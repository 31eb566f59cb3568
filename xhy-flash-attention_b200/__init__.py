"""xhy-flash-attention_b200 -- B200-native (sm_100a) drop-in for the fused-attention hot path of
starcrown001/xhy-flash-attention: the functions of ``flash_attn/flash_attn_interface.py`` over the C ABI of
``libflashattn_b200.so`` (include/fa_capi.h, include/fa_b200.h).

The directory name is not a Python identifier; import it with
``importlib.import_module("xhy-flash-attention_b200")`` (tests/conftest.py and bench.py do exactly that).
"""
__version__ = "0.1.0"

from .host_pipeline import HostPipeline  # noqa: F401
from . import bert_padding  # noqa: F401
from .varlen_shard import VarlenShard, flash_attn_varlen_sharded, lpt_assign, shard_varlen  # noqa: F401
from .flash_attn_interface import (  # noqa: F401
    flash_attn_fp8_func,
    flash_attn_func,
    flash_attn_kvpacked_func,
    flash_attn_qkvpacked_func,
    flash_attn_varlen_func,
    flash_attn_varlen_kvpacked_func,
    flash_attn_varlen_qkvpacked_func,
    flash_attn_with_kvcache,
    flashmask_attention,
)
