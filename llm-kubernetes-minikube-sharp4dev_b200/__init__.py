"""B200-native vector store + top-k search for the Minimal_RAG retrieval step.

The product is csrc/ (hand-written sm_100a CUDA behind the C ABI of include/ragstore.h).  This
Python package is the thin host side used by tests and bench.py: a ctypes binding (`native`) and a
mirror of the reference's RagIndex surface (`rag_index`).  There is no CPU scoring path here: if
libragstore.so or a CUDA device is missing, calls raise.
"""
from .native import (  # noqa: F401
    RAGSTORE_ADD_DEVICE, RAGSTORE_ADD_HOST, RAGSTORE_ADD_PRENORMALISED, RAGSTORE_BF16, RAGSTORE_F32,
    RAGSTORE_CREATE_BF16_SHADOW, RAGSTORE_CREATE_SIDECAR_OFF, RAGSTORE_CREATE_SIDECAR_ON,
    RAGSTORE_HIT_CERTIFIED, RAGSTORE_HIT_EXACT_ROWS, RagStore, RagStoreError, lib, lib_path,
)
from .rag_index import RagChunk, RagHit, RagIndex  # noqa: F401
from .sharded import ShardedSearcher, gather_layout, shard_range  # noqa: F401,E402
