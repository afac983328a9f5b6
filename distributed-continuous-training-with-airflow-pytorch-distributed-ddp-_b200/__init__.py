"""wxddp -- B200-native (sm_100a) implementation of the weather-MLP DDP training step and the
standardisation that feeds it (reference: jobs/train_lightning_ddp.py, jobs/preprocess.py).

The directory name carries hyphens (it is fixed by the project layout), so the importable alias is the
top-level `wxddp` package, which points its __path__ here.  Host side: Python + torch tensors as memory
owners; device side: libwxddp.so (hand-written CUDA behind the C ABI of include/wxddp.h).  No CPU fallback.
"""
from ._lib import (DROPOUT_MASK, DROPOUT_OFF, DROPOUT_PHILOX, IMPUTE_MEAN, IMPUTE_NONE, WX_F32, WX_F64,  # noqa: F401
                   WxError, lib_path, load)
from ._build import build_library  # noqa: F401
from .ops import Context, n_params  # noqa: F401

__version__ = "0.1.0"
