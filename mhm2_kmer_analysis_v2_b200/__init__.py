"""B200-native k-mer analysis (kcount) stage of MetaHipMer2 -- Python host layer over the C ABI.

Layout: csrc/ (sm_100a kernels + C ABI, built into libkcount_b200.so), _lib.py (ctypes binding), kcount.py (stage object
and the reference-shaped KmerDHT / analyze_kmers), exchange.py (NCCL supermer all-to-all), reads.py (seeded inputs).
"""
from ._lib import KcError, build  # noqa: F401
from .kcount import KcountStage, KmerDHT, KmerResult, PackedReads, analyze_kmers  # noqa: F401
from .reads import CONFIGS, PackedReadsLayout, PairSet, ReadSet, Synth, generate, merge_pairs  # noqa: F401
