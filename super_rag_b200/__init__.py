"""super_rag_b200 — B200-native hybrid-retrieval hot path for promoteAI/super-rag.

Exact cosine / inner-product top-k (tcgen05 GEMM fused with normalisation and top-k), BM25
posting-list scoring and RRF / weighted fusion as hand-written sm_100a CUDA kernels behind a
C ABI (include/srag_b200.h), plus the host-side mirror of the reference's connector / node
interfaces. Importing the package does not need a GPU; calling any op does.
"""
from ._lib import SragError, lib_path, load  # noqa: F401

__version__ = "0.1.0"
