"""basevarc_b200 -- B200-native (sm_100a) per-site `basetype` calling core behind a C ABI.

Only what the hot path needs lives here: `csrc/` (CUDA kernels + C-ABI library), the ctypes
mirror of the boundary (`_abi`, `caller`) and the packed-pileup tile (`pileup`).
"""
from ._abi import SITE_CORE_DTYPE, SITE_RESULT_DTYPE, Params, Tile, default_params, make_code  # noqa: F401
from .caller import (BaseTypeCaller, BvcError, CodedArrays, ShardedCaller, ShardGroup, load_library, measure_fp64_peak,  # noqa: F401
                     shard_unique_id)
from .pileup import CONFIGS, PileupTile, synth_config, synth_tile  # noqa: F401
