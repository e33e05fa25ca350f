"""B200-native (sm_100a) hot path behind the SGLang operator surface.

Scope: one AttentionBackend (paged decode / extend over ReqToTokenPool +
MHATokenToKVPool), FP8 and AWQ/GPTQ-INT4 LinearMethod.apply kernels, TP all-reduce at the
row-parallel boundaries.  All compute goes through the C-ABI library
`lib/libb200hotpath.so` (include/b200_hotpath.h); there is no CPU or eager fallback.
"""

__version__ = "0.1.0"
