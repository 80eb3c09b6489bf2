// NativeMethods.cs -- P/Invoke declarations for libragstore.so (include/ragstore.h).
// SOURCE ONLY: there is no .NET toolchain in the build environment, so this file has never been compiled
// or run; it shows the binding a maintainer of Minimal_RAG would add next to Helpers/RagIndex.cs.
using System;
using System.Runtime.InteropServices;

namespace Minimal_RAG.Helpers
{
    internal static partial class NativeMethods
    {
        const string Lib = "ragstore";   // resolves to libragstore.so on Linux

        public const int RAGSTORE_F32 = 0, RAGSTORE_BF16 = 1;
        public const uint RAGSTORE_ADD_HOST = 0;
        public const int RAGSTORE_HIT_CERTIFIED = 1, RAGSTORE_HIT_EXACT_ROWS = 2;
        public const uint RAGSTORE_CREATE_BF16_SHADOW = 4;   // fp32 store + bf16 image for the batched tensor-core nomination

        [LibraryImport(Lib)] public static partial int ragstore_create(int dim, int dtype, long capacityRows, int device, uint flags, out IntPtr store);
        // ONE handle over several GPUs of this process: same add / search / count calls (the singleton of Program.cs:35,43
        // keeps its shape when the corpus outgrows one GPU)
        [LibraryImport(Lib)] public static unsafe partial int ragstore_create_sharded(int dim, int dtype, long capacityRows, int* devices, int nDevices,
                                                                                     long blockRows, uint flags, out IntPtr store);
        [LibraryImport(Lib)] public static partial int ragstore_destroy(IntPtr store);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_add(IntPtr store, float* rows, long nRows, int rowDim, uint flags, out long firstRow);
        [LibraryImport(Lib)] public static partial int ragstore_count(IntPtr store, out long rows);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_search(IntPtr store, float* queries, int nQueries, int queryDim, int topK,
                                                                             long* outIdx, double* outScore, int* outFound, int* outFlags);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_last_error(byte* buf, int len);

        // one query per call, coalesced with concurrent callers into one corpus sweep (SURVEY 8(f) N1)
        [LibraryImport(Lib)] public static unsafe partial int ragstore_search_coalesced(IntPtr store, float* query, int queryDim, int topK,
                                                                                       long* outIdx, double* outScore, int* outFound, int* outFlags);
        [LibraryImport(Lib)] public static partial int ragstore_set_coalescing(IntPtr store, int windowUs, int maxBatch);
        // Program.cs:77-89,125-128 in one call: citation mask, best score, low-score flag (N2)
        [LibraryImport(Lib)] public static unsafe partial int ragstore_hit_filters(double* scores, int found, double evidenceMinScore, double bestFactor,
                                                                                  double lowScoreFloor, byte* outCited, out double best, out int lowScore);
        // capacity planning and scatter-gather appends (RagIndex.cs:8,48: List<T>.Capacity / AddRange)
        [LibraryImport(Lib)] public static partial int ragstore_reserve_rows(IntPtr store, long nRows);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_add_many(IntPtr store, float** blocks, long* blockRows, int nBlocks, int rowDim,
                                                                               uint flags, out long firstRow);
        // store image on disk (N3)
        [LibraryImport(Lib, StringMarshalling = StringMarshalling.Utf8)] public static partial int ragstore_save(IntPtr store, string path);
        [LibraryImport(Lib, StringMarshalling = StringMarshalling.Utf8)] public static partial int ragstore_load(string path, int device, long capacityRows, out IntPtr store);
        // row-sharded hosts (one process per GPU): per-shard hits, then the peer-memory exchange + merge in one launch
        [LibraryImport(Lib)] public static partial int ragstore_set_row_base(IntPtr store, long rowBase);
        [LibraryImport(Lib)] public static partial int ragstore_search_shard_device(IntPtr store, IntPtr dQuery, int topK, IntPtr dOut, IntPtr dFlags, IntPtr stream);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_search_shard_hostq(IntPtr store, float* hostQuery, int topK, IntPtr dOut, IntPtr dFlags, IntPtr stream);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_exchange_create(int device, int world, int rank, int maxRecords, int maxQueries,
                                                                                      out IntPtr exchange, byte* handleOut /* 64 bytes */);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_exchange_connect(IntPtr exchange, byte* allHandles /* world x 64 bytes, rank order */);
        [LibraryImport(Lib)] public static partial int ragstore_exchange_merge(IntPtr exchange, IntPtr dMine, int nQueries, int perShard, int topK,
                                                                              IntPtr dIdx, IntPtr dScore, IntPtr dFound, IntPtr dStatus, IntPtr stream);
        [LibraryImport(Lib)] public static partial int ragstore_exchange_merge_flags(IntPtr exchange, IntPtr dMine, IntPtr dMyFlags, int nQueries, int perShard, int topK,
                                                                                    IntPtr dIdx, IntPtr dScore, IntPtr dFound, IntPtr dFlags, IntPtr dStatus, IntPtr stream);
        // shard scan + exchange + merge in ONE launch per rank (give exactly one of dQuery / hostQuery)
        [LibraryImport(Lib)] public static unsafe partial int ragstore_search_shard_exchange(IntPtr store, IntPtr exchange, IntPtr dQuery, float* hostQuery, int topK,
                                                                                            IntPtr dMine, IntPtr dIdx, IntPtr dScore, IntPtr dFound, IntPtr dFlags,
                                                                                            IntPtr dStatus, IntPtr stream);
        [LibraryImport(Lib)] public static partial int ragstore_exchange_destroy(IntPtr exchange);

        public static unsafe void Check(int rc)
        {
            if (rc == 0) return;
            var buf = stackalloc byte[512];
            int n = ragstore_last_error(buf, 512);
            throw new InvalidOperationException($"ragstore error {rc}: {System.Text.Encoding.UTF8.GetString(buf, n)}");
        }
    }
}
