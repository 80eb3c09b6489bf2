"""Host-side mirror of the reference's RagIndex surface for the retrieval path.

Reference (Minimal_Agent_RAG/Minimal_RAG): Helpers/RagIndex.cs:6-11,48,59-67 and the records at
Program.cs:326-327.  Same member names, argument meaning and corner cases; the `float[] Embedding`
goes to the device store instead of the managed heap, the Id/Source/Text stay on the host in
insertion order (row id == list position, as in the reference).  The text pipeline
(BuildFromFolderAsync's chunking) and the embedding HTTP client are out of scope: embeddings come
from an injected `embedder(model, text) -> sequence of float` (Embedder.EmbedAsync's contract).
"""
from __future__ import annotations

import asyncio
import inspect
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import numpy as np

from .native import RAGSTORE_F32, RagStore


@dataclass(frozen=True)
class RagChunk:            # Program.cs:326
    Id: str
    Source: str
    Text: str
    Embedding: Optional[Sequence[float]] = None   # kept only until it is uploaded


@dataclass(frozen=True)
class RagHit:              # Program.cs:327
    Id: str
    Source: str
    Text: str
    Score: float           # C# double


class RagIndex:
    """Drop-in for `class RagIndex` (RagIndex.cs:6): store + brute-force cosine top-k."""

    def __init__(self, embedModel: str, embedder: Optional[Callable] = None, *, dim: Optional[int] = None,
                 capacity_rows: int = 1 << 20, dtype: int = RAGSTORE_F32, device: int = 0, coalesce: bool = False,
                 batch_embedder: Optional[Callable] = None, embed_batch: int = 64, devices: Optional[Sequence[int]] = None,
                 block_rows: int = 0, store_flags: int = 0):
        self._embedModel = embedModel                       # RagIndex.cs:9,11
        self._embedder = embedder
        self._meta: List[tuple] = []                        # (Id, Source, Text) per row, insertion order
        self._store: Optional[RagStore] = None
        self._dim, self._cap, self._dtype, self._device = dim, capacity_rows, dtype, device
        # `devices`: the same index spread over several GPUs of this process behind ONE handle (ragstore_create_sharded);
        # the object keeps its shape -- one RagIndex, as the reference's DI singleton (Program.cs:35,43)
        self._devices, self._block_rows, self._store_flags = (list(devices) if devices else None), block_rows, store_flags
        self._coalesce = coalesce          # concurrent QueryAsync callers share one corpus sweep
        # `batch_embedder(model, [texts]) -> [embeddings]` -- the array form of the embed endpoint that
        # Embedder.cs:19 already sends for one text -- turns one HTTP round trip per chunk (RagIndex.cs:42-50) into one
        # per `embed_batch` chunks; the rows then reach the device as one ragstore_add batch either way.
        self._batch_embedder, self._embed_batch = batch_embedder, max(1, int(embed_batch))
        self.last_flags = None

    # -- store side ---------------------------------------------------------------------------
    def _ensure(self, dim: int):
        if self._store is None:
            self._dim = self._dim or dim
            if self._devices and len(self._devices) > 1:
                self._store = RagStore.create_sharded(self._dim, self._cap, self._devices, dtype=self._dtype, block_rows=self._block_rows,
                                                      flags=self._store_flags)
            else:
                self._store = RagStore(self._dim, self._cap, self._dtype, self._devices[0] if self._devices else self._device, self._store_flags)

    def Add(self, chunk: RagChunk):
        """`_chunks.Add(new RagChunk(..))` (RagIndex.cs:48)."""
        self.AddRange([chunk])

    def AddRange(self, chunks: Sequence[RagChunk]):
        """Batched form of the store call: one upload/normalise launch for all embeddings."""
        if not chunks:
            return
        emb = np.asarray([c.Embedding for c in chunks], dtype=np.float32)
        self._ensure(emb.shape[1])
        first = self._store.add(emb)
        assert first == len(self._meta)
        self._meta.extend((c.Id, c.Source, c.Text) for c in chunks)

    @property
    def Count(self) -> int:
        return len(self._meta)

    async def BuildFromChunksAsync(self, items, logger=None):
        """The tail of BuildFromFolderAsync (RagIndex.cs:42-50): embed each cleaned chunk and store it.
        `items` yields (file_name, source_path, clean_text); ids follow `"{file}#{i}"` per file."""
        counters, batch, pending = {}, [], []

        async def flush():
            if not pending:
                return
            r = self._batch_embedder(self._embedModel, [t for _, _, t in pending])
            if inspect.isawaitable(r):
                r = await r
            if len(r) != len(pending):
                raise RuntimeError("Embedding batch returned a different number of vectors")   # cf. Embedder.cs:28
            for (cid, source, text), emb in zip(pending, r):
                batch.append(RagChunk(cid, source, text, np.asarray(emb, dtype=np.float32)))
            pending.clear()

        for file_name, source, text in items:
            if text is None or not str(text).strip():
                continue                                      # :45
            i = counters.get(file_name, 0)
            counters[file_name] = i + 1
            cid = f"{file_name}#{i}"
            if self._batch_embedder is not None:
                pending.append((cid, source, text))
                if len(pending) >= self._embed_batch:
                    await flush()
            else:
                emb = await self._embed(text)                 # :47
                batch.append(RagChunk(cid, source, text, emb))   # :48
        if self._batch_embedder is not None:
            await flush()
        self.AddRange(batch)
        if logger is not None:
            logger.info("[RAG] Totale chunks indicizzati: %d", self.Count)   # :55

    async def _embed(self, text: str):
        if self._embedder is None:
            raise RuntimeError("no embedder configured (Embedder.EmbedAsync is outside this library)")
        r = self._embedder(self._embedModel, text)
        if inspect.isawaitable(r):
            r = await r
        return np.asarray(r, dtype=np.float32)

    # -- search side --------------------------------------------------------------------------
    def QueryEmbedding(self, q: Sequence[float], topK: int = 5) -> List[RagHit]:
        """QueryAsync after the embedding call (RagIndex.cs:62-66)."""
        if self._store is None or not self._meta:
            return []                                         # empty index -> empty list
        if self._coalesce:
            idx1, score1, flags1 = self._store.search_coalesced(np.asarray(q, dtype=np.float32), topK)
            self.last_flags = int(flags1)
        else:
            idx, score, found, flags = self._store.search(np.asarray(q, dtype=np.float32), topK)
            idx1, score1 = idx[0, :int(found[0])], score[0, :int(found[0])]
            self.last_flags = int(flags[0])
        hits = []
        for i, sc in zip(idx1, score1):
            Id, Source, Text = self._meta[int(i)]
            hits.append(RagHit(Id, Source, Text, float(sc)))
        return hits

    @staticmethod
    def Citations(hits: Sequence[RagHit], evidenceMinScore: float = 0.35, bestFactor: float = 0.6) -> List[str]:
        """The /agent_rag citation rule (Program.cs:125-128): ids of hits with Score >= max(0.35, 0.6 * best)."""
        from .native import hit_filters
        if not hits:
            return []
        cited, _, _ = hit_filters([h.Score for h in hits], evidenceMinScore, bestFactor)
        return [h.Id for h, c in zip(hits, cited) if c]

    async def QueryAsync(self, query: str, topK: int = 5) -> List[RagHit]:
        """`Task<List<RagHit>> QueryAsync(string query, int topK = 5)` (RagIndex.cs:59)."""
        q = await self._embed(query)                          # :61
        return self.QueryEmbedding(q, topK)

    def Query(self, query: str, topK: int = 5) -> List[RagHit]:
        return asyncio.run(self.QueryAsync(query, topK))

    def close(self):
        if self._store is not None:
            self._store.close()
            self._store = None
