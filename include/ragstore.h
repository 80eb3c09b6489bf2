/*
 * ragstore.h -- C ABI of the B200-native vector store + top-k search that replaces the
 * in-memory RAG index of johnnypax/llm-kubernetes-minikube-sharp4dev.
 *
 * The reference has no FFI at all: its store is `List<RagChunk>` on the managed heap and its
 * search is a LINQ chain.  Every entry point below therefore cites the managed member whose
 * work it takes over (paths relative to Minimal_Agent_RAG/Minimal_RAG/ in the reference):
 *
 *   Helpers/RagIndex.cs:8        readonly List<RagChunk> _chunks          -> the store handle
 *   Helpers/RagIndex.cs:48       _chunks.Add(new RagChunk(.., emb))       -> ragstore_add
 *   Helpers/RagIndex.cs:59-67    QueryAsync: Select(Cosine).OrderByDescending.Take(max(1,k))
 *                                                                         -> ragstore_search
 *   Helpers/RagIndex.cs:124-135  static double Cosine(float[], float[])   -> the scoring kernels
 *   Program.cs:326-327           RagChunk(float[] Embedding) / RagHit(double Score)
 *                                                                         -> fp32 in, fp64 score out
 *
 * Threads: CUDA's current device is per thread.  Every call switches to the store's device and back when the calling
 * thread is on another one; call from a thread already on the store's device to avoid that cost (it is measurable:
 * DESIGN.md 4b).
 * Conventions: plain C types only; every call returns an int status (0 = RAGSTORE_OK) and never
 * unwinds across the boundary; the caller owns every in/out buffer; the library owns device
 * memory; handles are opaque.  INTEGRATION.md shows the P/Invoke declarations a maintainer
 * of the reference would add.
 */
#ifndef RAGSTORE_H_
#define RAGSTORE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define RAGSTORE_API __attribute__((visibility("default")))
#else
#define RAGSTORE_API
#endif

typedef struct ragstore ragstore_t;

/* status codes */
enum {
    RAGSTORE_OK = 0,
    RAGSTORE_ERR_INVALID = 1,      /* bad argument (null pointer, dim <= 0, wrong row_dim on add, ...) */
    RAGSTORE_ERR_CUDA = 2,         /* a CUDA call failed; ragstore_last_error has the text */
    RAGSTORE_ERR_CAPACITY = 3,     /* add would exceed the capacity fixed at create */
    RAGSTORE_ERR_UNSUPPORTED = 4,  /* dim / k outside what the kernels cover */
    RAGSTORE_ERR_NO_DEVICE = 5     /* no usable sm_100 device: there is NO CPU fallback */
};

/* storage element type of the device matrix (north_star (1): fp32, optionally bf16) */
enum { RAGSTORE_F32 = 0, RAGSTORE_BF16 = 1 };

/* flags for ragstore_create.  A row whose stored image differs from the floats that were added
 * (normalising changed it, or bf16 rounding did) can keep its original fp32 values in a sidecar that
 * only the rescoring stage reads, so returned scores are the reference's values for the row as added.
 * Default: sidecar on for fp32 stores, off for bf16 stores. */
enum {
    RAGSTORE_CREATE_DEFAULT = 0,
    RAGSTORE_CREATE_SIDECAR_ON = 1,
    RAGSTORE_CREATE_SIDECAR_OFF = 2,
    /* fp32 stores only (rows a multiple of 8 elements): the upload kernel also writes a bf16 image of every stored row
     * (+50 % device memory).  Batches of queries then NOMINATE their candidates from that image on the bf16 tensor-core
     * path -- half the bytes and half the tensor-core work of tf32 -- from a wider candidate set (128 per query), and
     * the returned hits are still rescored from the fp32 rows with the reference's arithmetic, so results do not change.
     * Single queries through ragstore_search scan the image as well (half the bytes per query: 10M x 768 in 2.3 ms instead
     * of 4.5 ms) on corpora large enough for that to pay; whatever the coarser scan cannot certify is redone on the fp32
     * rows.  The device-only entry points keep scanning the fp32 rows (they cannot retry). */
    RAGSTORE_CREATE_BF16_SHADOW = 4
};

/* flags for ragstore_add */
enum {
    RAGSTORE_ADD_HOST = 0,           /* `rows` is host memory (pageable or pinned) */
    RAGSTORE_ADD_DEVICE = 1,         /* `rows` is device memory on the store's device */
    RAGSTORE_ADD_PRENORMALISED = 2   /* store the rows as given (caller asserts unit norm); the
                                        per-row reciprocal norm is still measured and applied */
};

/* per-query result flags written by the search calls */
enum {
    RAGSTORE_HIT_CERTIFIED = 1,  /* the candidate margin proved that no row outside the rescored
                                    set can belong to the top-k (DESIGN.md "certificate") */
    RAGSTORE_HIT_EXACT_ROWS = 2  /* every returned row was rescored from the exact fp32 values that
                                    were added (stored bit-for-bit, or kept in the sidecar), so its
                                    fp64 score is the reference's value, not a 1e-7 neighbour */
};

/* one rescored hit as exchanged between shards: 16 bytes, (score desc, id asc) is the total order */
typedef struct ragstore_cand {
    double score;   /* fp64 cosine, arithmetic of RagIndex.cs:124-135 */
    int64_t id;     /* global row id = row_base + local row; -1 marks an empty slot */
} ragstore_cand;

/* ---- life cycle -------------------------------------------------------------------------- */

/* Replaces `new RagIndex(..)` + `List<RagChunk> _chunks = new()` (RagIndex.cs:8,11).
 * dim: embedding length (the reference takes whatever the model returns; fixed here at create).
 * capacity_rows: upper bound on rows; only ADDRESS SPACE is reserved for it, physical HBM is mapped
 *   as rows arrive (append-only, like the reference's list), so a generous bound is free.
 * device: CUDA ordinal.  Fails with RAGSTORE_ERR_NO_DEVICE when no GPU is present. */
RAGSTORE_API int ragstore_create(int dim, int dtype, int64_t capacity_rows, int device,
                                 uint32_t flags, ragstore_t** out);

/* The same store over several GPUs of THIS process, behind the same handle.  The reference's index is one object in one
 * process -- a DI singleton (Program.cs:35,43) that every handler queries (Program.cs:72,116) -- so a corpus larger than
 * one GPU (100M x 768) must still be one `RagIndex` holding one handle.  ragstore_add deals the rows to the
 * devices in blocks of block_rows (0 = 65536), round-robin, in insertion order, and a row's id stays its place in that
 * order (RagIndex.cs:48); ragstore_search launches every shard's scan from the calling thread on per-device streams,
 * the shards store their rescored hits straight into the first device's memory (peer access over NVLink; the copy
 * engine where there is none) and one merge kernel there returns the hits in the order a single index would
 * (RagIndex.cs:64-65).  ragstore_add, ragstore_add_many, ragstore_reserve_rows, ragstore_count, ragstore_search (any
 * top_k up to 8128, batches on the tensor cores of every device), ragstore_search_coalesced, ragstore_info and
 * ragstore_destroy accept the handle; the per-device entry points (ragstore_search_device, ragstore_search_shard_*,
 * ragstore_device_views, ragstore_set_row_*) return RAGSTORE_ERR_INVALID.  ragstore_save writes all shards into one file;
 * ragstore_load_sharded maps it back over the same NUMBER of devices (rows are dealt by shard count).
 * A device may be listed more than once (several shards on one GPU). */
RAGSTORE_API int ragstore_create_sharded(int dim, int dtype, int64_t capacity_rows, const int* devices, int n_devices,
                                         int64_t block_rows, uint32_t flags, ragstore_t** out);

RAGSTORE_API int ragstore_destroy(ragstore_t* s);

/* ---- store side --------------------------------------------------------------------------- */

/* Replaces `_chunks.Add(new RagChunk(id, file, text, emb))` (RagIndex.cs:48), batched: n_rows
 * embeddings, row-major, row_dim floats each.  Rows are L2-normalised by the upload kernel and
 * appended; the first new row id (== insertion order == list position, as in the reference) is
 * returned.  The reference accepts any length at insert and scores -1 at query time
 * (RagIndex.cs:126); a fixed-width matrix cannot, so row_dim != dim is RAGSTORE_ERR_INVALID.
 * Returns after the rows are visible to searches.  Calls are serialised internally and may
 * overlap with searches; a search sees a prefix of the appended rows. */
RAGSTORE_API int ragstore_add(ragstore_t* s, const float* rows, int64_t n_rows, int row_dim,
                              uint32_t flags, int64_t* out_first_row);

/* Scatter-gather form of ragstore_add: appends n_blocks host (or device, per `flags`) blocks of block_rows[b] rows each,
 * in order, exactly as n_blocks calls would -- every block becomes visible to searches as soon as it is in, and
 * out_first_row receives the row id of the first row of block 0.  For hosts that queue up embedding responses
 * (RagIndex.cs:42-50 produces one row at a time) and want one native call per flush. */
RAGSTORE_API int ragstore_add_many(ragstore_t* s, const float* const* blocks, const int64_t* block_rows, int n_blocks,
                                   int row_dim, uint32_t flags, int64_t* out_first_row);

/* Maps device memory for the first n_rows rows now instead of as they arrive (like List<T>.Capacity for the reference's
 * `_chunks`, RagIndex.cs:8).  Optional: appends map memory on demand, in geometrically growing steps; a server that
 * knows its corpus size calls this once at start-up so that no append ever pays for page-table work (which is a
 * device-wide operation and, on hosts with several GPUs in peer-to-peer contact, stalls running searches too). */
RAGSTORE_API int ragstore_reserve_rows(ragstore_t* s, int64_t n_rows);

/* Replaces `_chunks.Count` (RagIndex.cs:55). */
RAGSTORE_API int ragstore_count(ragstore_t* s, int64_t* out_rows);

/* Global id of local row 0 (row-sharded corpora).  Default 0. */
RAGSTORE_API int ragstore_set_row_base(ragstore_t* s, int64_t row_base);

/* Streaming ingest over several shards: batches of block_rows rows are dealt
 * round-robin to n_shards stores; with this set, local row r of shard `shard_index` reports the global id
 * ((r / block_rows) * n_shards + shard_index) * block_rows + r % block_rows (+ row_base), i.e. its place in
 * the global insertion order -- the row id the reference's single list would have given it (RagIndex.cs:48).
 * block_rows = 0 switches back to contiguous ids. */
RAGSTORE_API int ragstore_set_row_blocking(ragstore_t* s, int64_t block_rows, int n_shards, int shard_index);

/* ---- search side -------------------------------------------------------------------------- */

/* Replaces the body of QueryAsync after the embedding call (RagIndex.cs:62-66) for n_queries
 * queries at once: host buffers in, host buffers out, synchronous, re-entrant.
 *   queries   n_queries x query_dim fp32, row-major (raw, not normalised)
 *   top_k     as passed to QueryAsync; max(1, top_k) rows are requested (RagIndex.cs:65)
 *   out_idx / out_score   n_queries x max(1,top_k); row ids best first, fp64 scores
 *   out_found per query: min(count, max(1,top_k))    out_flags per query: RAGSTORE_HIT_* (may be NULL)
 * Mirrors the reference's corner cases: empty store -> 0 hits; query_dim != dim -> every score is
 * -1, so rows 0..found-1 in insertion order with score -1.0 (RagIndex.cs:126 + stable sort).
 * Any top_k is accepted; up to 256 hits come from one kernel launch per query, more (up to 8128) from
 * repeated scans over disjoint candidate ranges.  More than 8128 hits: RAGSTORE_ERR_UNSUPPORTED. */
RAGSTORE_API int ragstore_search(ragstore_t* s, const float* queries, int n_queries, int query_dim,
                                 int top_k, int64_t* out_idx, double* out_score,
                                 int32_t* out_found, int32_t* out_flags);

/* the reference serves /rag/search and /agent_rag concurrently from one singleton index
 * (Program.cs:43,62,106), each request scanning the whole list on its own thread-pool thread.  This entry
 * point takes ONE query per call, like QueryAsync, but callers that arrive within a short window are
 * answered by one sweep of the corpus (the batched tensor-core path, from 2-3 callers up).  Same results as
 * ragstore_search; thread-safe; blocks until this caller's hits are ready. */
RAGSTORE_API int ragstore_search_coalesced(ragstore_t* s, const float* query, int query_dim, int top_k,
                                           int64_t* out_idx, double* out_score, int32_t* out_found,
                                           int32_t* out_flags);
/* window_us: how long every batch's first caller waits for company (default 0: nobody waits; callers that arrive while
 * a batch is executing form the next batch by themselves).  max_batch: queries per sweep (default 256). */
RAGSTORE_API int ragstore_set_coalescing(ragstore_t* s, int window_us, int max_batch);

/* Same work with every buffer already in device memory and no host synchronisation: enqueues on
 * `stream` (a cudaStream_t; NULL = the store's own stream) and returns.  query_dim must equal dim; at most
 * 256 hits per query (min(count, max(1,top_k)) <= 256), because nothing can be read back to decide on more passes.
 * Not re-entrant: one in-flight device search per store (it uses the store's scratch). */
RAGSTORE_API int ragstore_search_device(ragstore_t* s, const float* d_queries, int n_queries,
                                        int top_k, int64_t* d_idx, double* d_score,
                                        int32_t* d_found, int32_t* d_flags, void* stream);

/* Row-sharded search, step 1 (per shard): the shard's own rescored top max(1,top_k) as
 * `ragstore_cand` records (global ids), best first, padded with {-inf, -1}.  d_out has
 * max(1,top_k) records.  The records of all shards are then all-gathered (NCCL) by the host. */
RAGSTORE_API int ragstore_search_shard_device(ragstore_t* s, const float* d_query, int top_k,
                                              ragstore_cand* d_out, int32_t* d_flags, void* stream);

/* Step 1 with the query still in HOST memory (any host memory; read before the call returns): it travels inside the
 * kernel's parameter block, so a rank that receives the query from its host process needs no host->device copy in
 * front of the launch.  dim <= 1024.  With the outputs of ragstore_exchange_merge pointed at page-locked host memory
 * a sharded query is two launches and one stream wait per rank, end to end. */
RAGSTORE_API int ragstore_search_shard_hostq(ragstore_t* s, const float* h_query, int top_k,
                                             ragstore_cand* d_out, int32_t* d_flags, void* stream);

/* Batched form of step 1: n_queries queries at once (tensor-core path when eligible); d_out holds
 * n_queries x max(1,top_k) records, query-major. */
RAGSTORE_API int ragstore_search_shard_batch_device(ragstore_t* s, const float* d_queries, int n_queries,
                                                    int top_k, ragstore_cand* d_out, int32_t* d_flags,
                                                    void* stream);

/* Batched form of step 2: d_gathered is [n_shards][n_queries][per_shard] (what an all-gather of the
 * per-rank buffers of the call above produces); outputs are n_queries x max(1,top_k). */
RAGSTORE_API int ragstore_merge_shards_batch_device(const ragstore_cand* d_gathered, int n_shards,
                                                    int n_queries, int per_shard, int top_k,
                                                    int64_t* d_idx, double* d_score, int32_t* d_found,
                                                    int device, void* stream);

/* Row-sharded search, step 2 (every rank, after the all-gather): merge n_shards x per_shard
 * records into the global top max(1,top_k) under (score desc, id asc) -- the order a single
 * RagIndex over the concatenated corpus would return (RagIndex.cs:64-65). */
RAGSTORE_API int ragstore_merge_shards_device(const ragstore_cand* d_gathered, int n_shards,
                                              int per_shard, int top_k, int64_t* d_idx,
                                              double* d_score, int32_t* d_found, int device,
                                              void* stream);

/* Row-sharded search, steps "all-gather" and 2 fused into one launch over peer memory (one process per GPU
 * on one node, GPUs with peer access -- NVLink / NVSwitch).  Every rank owns an INBOX; a call stores this
 * rank's records into every rank's inbox, waits for everybody else's, and merges -- no NCCL call and no
 * host round trip on the query path.  Replaces, for the sharded form of RagIndex.QueryAsync
 * (RagIndex.cs:59-67), the host-side all-gather between ragstore_search_shard*_device and
 * ragstore_merge_shards*_device; results are identical.
 *   create   allocates the inbox for at most max_queries queries and max_records records
 *            (n_queries * per_shard) per call and writes the RAGSTORE_EXCHANGE_HANDLE_BYTES-byte handle
 *            that peers need; the caller moves the handles between the processes (any transport).
 *   connect  all_handles = world handles in rank order (own one included); maps the peers' inboxes.
 *   merge    d_mine = this rank's n_queries x per_shard records (what ragstore_search_shard*_device
 *            wrote); outputs as ragstore_merge_shards_batch_device; d_status (n_queries, may be null)
 *            receives 0, or 1 when a peer's records did not arrive within 5 s (results then invalid).
 *            Collective: every rank must make the same sequence of calls with the same n_queries and
 *            per_shard, each rank on one stream of its own. */
#define RAGSTORE_EXCHANGE_HANDLE_BYTES 64
typedef struct ragstore_exchange ragstore_exchange_t;
RAGSTORE_API int ragstore_exchange_create(int device, int world, int rank, int max_records, int max_queries,
                                          ragstore_exchange_t** out, void* handle_out);
RAGSTORE_API int ragstore_exchange_connect(ragstore_exchange_t* x, const void* all_handles);
RAGSTORE_API int ragstore_exchange_merge(ragstore_exchange_t* x, const ragstore_cand* d_mine, int n_queries,
                                         int per_shard, int top_k, int64_t* d_idx, double* d_score,
                                         int32_t* d_found, int32_t* d_status, void* stream);
/* The same with the certificate: d_my_flags (n_queries, may be null = both bits) are this rank's RAGSTORE_HIT_* bits as
 * written by ragstore_search_shard*_device; they travel with the records, and d_flags (n_queries, may be null) receives
 * the AND over all ranks -- a merged result is certified / exact only if every shard's part of it is. */
RAGSTORE_API int ragstore_exchange_merge_flags(ragstore_exchange_t* x, const ragstore_cand* d_mine, const int32_t* d_my_flags,
                                               int n_queries, int per_shard, int top_k, int64_t* d_idx, double* d_score,
                                               int32_t* d_found, int32_t* d_flags, int32_t* d_status, void* stream);
/* Steps 1, "all-gather" and 2 of the row-sharded search in ONE launch per rank: the shard's scan kernel, whose last CTA
 * -- the one that rescored and ranked the shard's hits -- also stores them into every rank's inbox, waits for the other
 * ranks' and ranks the G x k records.  Replaces, per rank, the whole of the sharded RagIndex.QueryAsync
 * (RagIndex.cs:59-67).  Exactly one of d_query (device) / h_query (host memory, read before the call returns; it rides
 * in the kernel parameters, dim <= 1024) is given.  d_mine: max(1,top_k) records of scratch on this device (receives
 * this shard's own hits); d_idx / d_score: max(1,top_k) merged hits (device or page-locked host memory); d_flags: the
 * RAGSTORE_HIT_* bits ANDed over all ranks; d_status as ragstore_exchange_merge.  Collective like
 * ragstore_exchange_merge (same sequence of calls on every rank); top_k <= 256. */
RAGSTORE_API int ragstore_search_shard_exchange(ragstore_t* s, ragstore_exchange_t* x, const float* d_query, const float* h_query,
                                                int top_k, ragstore_cand* d_mine, int64_t* d_idx, double* d_score,
                                                int32_t* d_found, int32_t* d_flags, int32_t* d_status, void* stream);
RAGSTORE_API int ragstore_exchange_destroy(ragstore_exchange_t* x);

/* No counterpart in the reference: an extra search entry point for hosts that are not allowed to
 * allocate device memory -- a server that gave all of HBM to its own arena at start-up, or a request
 * thread that cannot afford the device-wide synchronisation and driver locks of cudaMalloc/cudaFree.
 * It runs the same scan + top-k + exact-rescoring launch as ragstore_search_device on buffers the
 * caller owns, and calls no cudaMalloc, cudaFree or other allocating routine, ever.
 *   d_matrix  n_rows x dim, row-major, fp32 or bf16 (dtype), rows a multiple of 16 bytes, unit norm (RAGSTORE_HIT_CERTIFIED
 *             is computed for rows of norm >= 1/2; shorter rows are still rescored exactly, but the flag may over-promise)
 *   d_rnorm   reciprocal norm of each stored row, fp32, readable up to the next multiple of 4 rows
 *   d_query   dim raw fp32 values;  outputs as ragstore_search_device for one query
 * The kernel's scratch (per-CTA candidate lists + a ticket) lives in a fixed array inside the
 * library image: ragstore_noalloc_scratch_bytes() (about 321 KB) is held on every GPU that loads the
 * library, until the process exits.  Limits that follow from the fixed size: top_k <= 256 and at most
 * 160 SMs; beyond them the call returns RAGSTORE_ERR_UNSUPPORTED and says why.
 * The first call on each device resolves the array's device address, and -- CUDA loads modules lazily by default -- may
 * make the driver load the kernel for that dtype / row width together with the array itself; make one call per device
 * and per matrix shape at start-up, before any stream capture begins and before the arena takes the last of HBM.  All calls share the one array: it keeps the previous call's contents (the
 * call clears the ticket it needs zeroed, in stream order), and two calls whose kernels overlap on the
 * same GPU race on it -- issue them on one stream, or otherwise order them.
 * Errors of this entry point are reported by ragstore_noalloc_last_error(). */
RAGSTORE_API int ragstore_search_noalloc(const void* d_matrix, const float* d_rnorm, int64_t n_rows,
                                         int dim, int dtype, const float* d_query, int top_k,
                                         int64_t* d_idx, double* d_score, int32_t* d_found,
                                         int32_t* d_flags, int device, void* stream);
RAGSTORE_API int64_t ragstore_noalloc_scratch_bytes(void);
RAGSTORE_API const char* ragstore_noalloc_last_error(void);

/* the score post-filters the two callers apply to a hit list, computed in one place:
 *   out_best       hits.Max(h => h.Score)                                  (Program.cs:77, :125)
 *   out_low_score  best < low_score_floor   (0.20 in /rag/search)           (Program.cs:78,82)
 *   out_cited[i]   scores[i] >= max(evidence_min_score, best * best_factor) (0.35 and 0.6 in /agent_rag, :126)
 * Pure host arithmetic on <= k doubles; any output may be NULL. */
RAGSTORE_API int ragstore_hit_filters(const double* scores, int found, double evidence_min_score,
                                      double best_factor, double low_score_floor, uint8_t* out_cited,
                                      double* out_best, int32_t* out_low_score);

/* the reference rebuilds its index from text at every start (Program.cs:50: one embedding
 * HTTP call per chunk).  ragstore_save writes the device-resident store (normalised matrix, reciprocal
 * norms, exact flags, sidecar) as one flat little-endian image; ragstore_load maps it back into HBM.
 * A loaded store answers exactly as the saved one did.  capacity_rows: 0 = just the saved rows. */
RAGSTORE_API int ragstore_save(ragstore_t* s, const char* path);
RAGSTORE_API int ragstore_load(const char* path, int device, int64_t capacity_rows, ragstore_t** out);
/* The image of a multi-device handle (ragstore_save on a handle from ragstore_create_sharded) back over n_devices GPUs;
 * n_devices must equal the number of shards it was saved with.  A store created with RAGSTORE_CREATE_BF16_SHADOW gets its
 * bf16 image back on load (rebuilt from the fp32 rows; it is not part of the file). */
RAGSTORE_API int ragstore_load_sharded(const char* path, const int* devices, int n_devices, int64_t capacity_rows, ragstore_t** out);

/* ---- utilities ----------------------------------------------------------------------------- */

/* Last error text of the calling thread (empty string if none). Returns the length written. */
RAGSTORE_API int ragstore_last_error(char* buf, int buf_len);

/* Library/ABI version (major*1000 + minor). */
RAGSTORE_API int ragstore_version(void);

/* The RAGSTORE_BATCH_* environment knobs (kernel-variant experiments) are read once, at first use; tests
 * that switch them inside one process call this afterwards. */
RAGSTORE_API int ragstore_debug_reload_env(void);

/* Store facts for tests and the bench: fills up to n of
 * [dim, ld (padded row length), dtype, count, capacity, n_inexact_rows, device, sm_count,
 *  has_sidecar, sidecar_rows, matrix_bytes_mapped, uses_vmm, scan_launches, batch_launches,
 *  last_add_kernel_ns (device time of the upload/normalise launches of the last ragstore_add),
 *  scratch_overlaps (scan launches that found their scratch in use by another launch: a caller broke the
 *  "one in-flight device search per store" rule of ragstore_search_device; results of such launches are garbage)]. */
RAGSTORE_API int ragstore_info(ragstore_t* s, int64_t* out, int n);

/* Profiling aid: %globaltimer (ns) at the phase boundaries of the store's most recent
 * ragstore_search_device launch: [start, scan done, lists merged, rows staged, rescored, end]. */
RAGSTORE_API int ragstore_debug_stamps(ragstore_t* s, uint64_t* out, int n);

/* Raw device views of the store (matrix, per-row reciprocal norms) for tests and benches. */
RAGSTORE_API int ragstore_device_views(ragstore_t* s, void** d_matrix, float** d_rnorm,
                                       uint8_t** d_exact);

/* Fills a device fp32 buffer with the synthetic benchmark corpus (SURVEY.md 8(d) "Synthetic
 * inputs"; bit-identical to oracle/ragindex_oracle.c:rag_synth_rows).  Not a reference function:
 * benchmark plumbing so that a 30 GB corpus never crosses PCIe. */
RAGSTORE_API int ragstore_synth_fill_device(float* d_out, int64_t row0, int64_t n_rows, int dim,
                                            int64_t ld, uint64_t seed, int round_bf16, int device,
                                            void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RAGSTORE_H_ */
