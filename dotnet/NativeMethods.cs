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

        [LibraryImport(Lib)] public static partial int ragstore_create(int dim, int dtype, long capacityRows, int device, uint flags, out IntPtr store);
        [LibraryImport(Lib)] public static partial int ragstore_destroy(IntPtr store);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_add(IntPtr store, float* rows, long nRows, int rowDim, uint flags, out long firstRow);
        [LibraryImport(Lib)] public static partial int ragstore_count(IntPtr store, out long rows);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_search(IntPtr store, float* queries, int nQueries, int queryDim, int topK,
                                                                             long* outIdx, double* outScore, int* outFound, int* outFlags);
        [LibraryImport(Lib)] public static unsafe partial int ragstore_last_error(byte* buf, int len);

        public static unsafe void Check(int rc)
        {
            if (rc == 0) return;
            var buf = stackalloc byte[512];
            int n = ragstore_last_error(buf, 512);
            throw new InvalidOperationException($"ragstore error {rc}: {System.Text.Encoding.UTF8.GetString(buf, n)}");
        }
    }
}
