"""volcanosv_b200 -- B200-native large-indel hot path of VolcanoSV-vc.

Python host (argv/file contracts of the reference operators) over a C-ABI CUDA library
(`libvolcano_b200.so`, include/volcanosv_b200.h).  No CPU fallback.
"""
from .lib import (Context, VsvError, SIG_DTYPE, MODE_CONTIG, MODE_READS, MODE_CUTE, DEL, INS,  # noqa: F401
                  SPLIT_HIFI, SPLIT_ONT, SPLIT_CLR, SPLIT_READS, load_library, LIB_PATH)

__version__ = "0.1.0"
