// RagIndex.cs -- device-backed drop-in for Minimal_RAG.Helpers.RagIndex (same public members, same results).
// SOURCE ONLY: never compiled here (no dotnet).  The text pipeline and the embedding client are the
// reference's own and stay where they are; only the store call and the search body change.
using System;
using System.Collections.Generic;
using System.Threading.Tasks;
using Microsoft.Extensions.Logging;

namespace Minimal_RAG.Helpers
{
    class RagIndex : IDisposable
    {
        // what used to be List<RagChunk>: the text stays managed, the embedding lives in HBM
        readonly List<(string Id, string Source, string Text)> _meta = new();
        readonly string _embedModel;
        IntPtr _store;
        int _dim;
        const long CapacityRows = 1L << 26;   // address space only; HBM is mapped as rows arrive
        // GPUs the index is spread over.  One entry: a single-device store.  Several: ONE handle over all of them
        // (ragstore_create_sharded) -- rows are dealt round-robin in blocks, ids stay insertion order, QueryAsync does not change.
        readonly int[] _devices;
        public bool Coalesce { get; set; } = true;

        public RagIndex(string embedModel) : this(embedModel, new[] { 0 }) { }
        public RagIndex(string embedModel, int[] devices) { _embedModel = embedModel; _devices = devices; }

        public async Task BuildFromFolderAsync(string folder, int chunkSize, int overlap, ILogger logger)
        {
            // ... unchanged enumeration / SplitByMarkdownHeaders / Sanitize from the reference ...
            // inside the per-chunk loop, the reference's
            //     _chunks.Add(new RagChunk($"{Path.GetFileName(file)}#{i++}", file, clean, emb));
            // becomes:
            //     Add($"{Path.GetFileName(file)}#{i++}", file, clean, emb);
            await Task.CompletedTask;
        }

        unsafe void Add(string id, string source, string text, float[] emb)
        {
            if (_store == IntPtr.Zero)
            {
                _dim = emb.Length;
                if (_devices.Length == 1)
                    NativeMethods.Check(NativeMethods.ragstore_create(_dim, NativeMethods.RAGSTORE_F32, CapacityRows, _devices[0], 0, out _store));
                else
                    fixed (int* pd = _devices)
                        NativeMethods.Check(NativeMethods.ragstore_create_sharded(_dim, NativeMethods.RAGSTORE_F32, CapacityRows, pd, _devices.Length,
                                                                                  0 /* 65536-row blocks */, 0, out _store));
            }
            fixed (float* p = emb)
                NativeMethods.Check(NativeMethods.ragstore_add(_store, p, 1, emb.Length, NativeMethods.RAGSTORE_ADD_HOST, out long row));
            _meta.Add((id, source, text));       // row id == list position, as in the reference
        }

        public async Task<List<RagHit>> QueryAsync(string query, int topK = 5)
        {
            var q = await Embedder.EmbedAsync(_embedModel, query);
            return Search(q, topK);
        }

        unsafe List<RagHit> Search(float[] q, int topK)
        {
            var hits = new List<RagHit>();
            if (_store == IntPtr.Zero) return hits;                     // empty index -> empty list
            int want = Math.Max(1, topK);
            var idx = new long[want]; var score = new double[want];
            int found, flags;
            // Under load the ASP.NET handlers (Program.cs:62,106) call QueryAsync concurrently: with Coalesce set, callers that
            // arrive together are answered by ONE sweep of the corpus on the tensor cores (same hits, same order).
            fixed (float* pq = q) fixed (long* pi = idx) fixed (double* ps = score)
                NativeMethods.Check(Coalesce
                    ? NativeMethods.ragstore_search_coalesced(_store, pq, q.Length, topK, pi, ps, &found, &flags)
                    : NativeMethods.ragstore_search(_store, pq, 1, q.Length, topK, pi, ps, &found, &flags));
            for (int j = 0; j < found; j++)
            {
                var m = _meta[(int)idx[j]];
                hits.Add(new RagHit(m.Id, m.Source, m.Text, score[j]));
            }
            return hits;
        }

        public void Dispose()
        {
            if (_store != IntPtr.Zero) { NativeMethods.ragstore_destroy(_store); _store = IntPtr.Zero; }
        }
    }
}
