"""ipu_path_trace_b200: B200-native (sm_100a) drop-in for the path-tracing + NIF hot path of
markp-gc/ipu_path_trace. The compute lives in libb200pt.so (hand-written CUDA behind the C ABI of
include/b200pt.h); this package is the thin host mirror of the reference's programs and loaders.
There is no CPU or PyTorch fallback: without the CUDA library every entry point raises."""
from ._lib import EXPORTS, LIB_PATH, build, load
from .load_balancer import (PAD_COORD, TRACE_RECORD, allocate_work_by_path_length, clear_accumulators,
                            create_work_list_for_image, pad_work_list, shard_tiles)
from .nif import NifMetaData, NifModel, load_assets, read_nif_bin, synthetic_model, write_nif_bin
from .tracer import (AA_NOISE_TYPES, AA_NORMAL, AA_TRUNCATED_NORMAL, AA_UNIFORM, NIF_AUTO, NIF_SIMT, NIF_TCGEN05,
                     B200PTError, Counters, PathTracer, PinnedRecords)

__all__ = [n for n in dir() if not n.startswith("_")]
