"""B200-native seeding + extension hot path of bwa-mem-scale.

The compute lives in `libbwagpu.so` (CUDA, sm_100a) behind the C ABI in include/bwagpu.h; this
package is the host-side mirror of the reference's interface for that path.
"""
from .lib import BwaGpuError, BswParams, Context, LIB_PATH, load  # noqa: F401
from .bsw import BandedPairWiseSW  # noqa: F401
from .seeding import FMI_search, SMEM_DTYPE, seed_opt, pack_reads  # noqa: F401
from .fmindex import FMIndex, build_fm_index  # noqa: F401
from .ksw import MateSW, KSWR_DTYPE, fill_scmat  # noqa: F401
from .extend import Extender, ext_opt, chain_opt, mem_core_device, CHAIN_DTYPE, SEED_DTYPE, REG_DTYPE  # noqa: F401
from .emf import PerfectTable, PerfectMatch, SEED_ENTRY_DTYPE, PERFECT_DTYPE  # noqa: F401
from .shard import shard_bounds, shard_reads, merge_smems, merge_regions  # noqa: F401
from .ert import ErtIndex, ErtSeeding, ERT_MEM_DTYPE, mem_core_ert_device  # noqa: F401
from .cigar import RegToAln, cigar_opt, CIGAR_JOB_DTYPE, CIGAR_OUT_DTYPE  # noqa: F401
from .mem import MemChunk, mem_opt, BATCH_SIZE  # noqa: F401
