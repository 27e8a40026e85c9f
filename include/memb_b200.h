/* memb_b200 — C ABI of the B200-native batched embedding lookup.
 *
 * This is the drop-in boundary for ONE path of thousandvoices/memb: what the
 * `_memb` pybind11 module (reference: python/memb_bindings.cpp:36-69) needs from
 * `memb::Reader` (reference: src/reader.h:11-40), re-expressed as plain C so any
 * FFI can bind it.  Every entry point cites the reference interface it replaces.
 *
 * Conventions
 *   - every function returns a memb_b200_status; on failure
 *     memb_b200_last_error() holds a thread-local message.  Lookups themselves
 *     never fail for unknown words: a missing word is a row of +0.0f
 *     (src/reader.cpp:41-47).
 *   - row id := index of the word in keys() (ascending byte order; the
 *     reference has no row-id API, it is the index lower_bound returns,
 *     src/trained_compression.cpp:115-127).  -1 (any negative or out-of-range id)
 *     means "missing" and decodes to zeros.
 *   - queries are byte strings [offsets[i], offsets[i+1]) of one blob; like the
 *     reference's strcmp they are cut at their first NUL byte.
 *   - buffers are owned by the caller; `*_kind` says where a pointer lives.
 *   - there is no CPU fallback: open() fails with MEMB_B200_ERR_NO_DEVICE when
 *     no sm_100 GPU is usable.
 */
#ifndef MEMB_B200_H
#define MEMB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct memb_b200_reader memb_b200_reader;
typedef struct memb_b200_builder memb_b200_builder;

typedef enum memb_b200_status {
    MEMB_B200_OK = 0,
    MEMB_B200_ERR_IO = 1,          /* file missing/unreadable (reference: boost mmap throws, src/tests.cpp:134-139) */
    MEMB_B200_ERR_FORMAT = 2,      /* "File format verification failed" & structural errors (src/reader.cpp:104-111) */
    MEMB_B200_ERR_UNSUPPORTED = 3, /* "Storage strategy %s is not supported" (src/compression_strategy.cpp:36-38) */
    MEMB_B200_ERR_CUDA = 4,
    MEMB_B200_ERR_ARGUMENT = 5,
    MEMB_B200_ERR_NO_DEVICE = 6,
    MEMB_B200_ERR_BUILDER = 7      /* dimension mismatch / duplicate word (src/builder.cpp:34-48) */
} memb_b200_status;

typedef enum memb_b200_mem_kind { MEMB_B200_HOST = 0, MEMB_B200_DEVICE = 1 } memb_b200_mem_kind;

/* wire::Storage, src/flatbuffers/embeddings.fbs:7-11 */
typedef enum memb_b200_storage { MEMB_B200_FULL = 1, MEMB_B200_UNIFORM = 2, MEMB_B200_TRAINED = 3 } memb_b200_storage;

/* python/memb/readers_union.py:35-38 */
typedef enum memb_b200_union_mode { MEMB_B200_AVERAGE = 0, MEMB_B200_CONCATENATE = 1 } memb_b200_union_mode;

typedef struct memb_b200_load_stats {
    double parse_ms;       /* mmap + flatbuffer walk + validation on the host */
    double upload_ms;      /* H2D of blob / offsets / vocabulary */
    double prepare_ms;     /* device-side row measuring + re-pack + hash build */
    uint64_t device_bytes; /* HBM held by this reader */
    uint64_t payload_bytes;/* sum of the rows' compressed bytes as stored in the file */
    uint32_t max_code_len; /* trained: longest prefix code; 0 otherwise */
    uint32_t lut_bits;     /* trained: direct-table width used on the device */
    uint32_t lut_replicas; /* trained: 32 = one LUT copy per shared-memory bank, 1 = single copy */
    uint32_t codebook_size;
} memb_b200_load_stats;

const char* memb_b200_last_error(void);

/* Number of usable sm_100 devices (0 when there is no GPU / wrong arch). */
int memb_b200_device_count(void);

/* Reader::Reader(filename, numThreads) — src/reader.cpp:23-29.  `device` is the
 * CUDA ordinal to load onto; num_threads (0 = all cores) only drives the host
 * parse, there is no CPU decode. */
memb_b200_status memb_b200_open(const char* path, int device, unsigned num_threads, memb_b200_reader** out);
void memb_b200_close(memb_b200_reader* reader);

/* Reader::dim() — src/reader.cpp:31-34 */
uint32_t memb_b200_dim(const memb_b200_reader* reader);
uint64_t memb_b200_rows(const memb_b200_reader* reader);
/* flatIndex_->storage_type() — src/reader.cpp:26 */
int memb_b200_storage_type(const memb_b200_reader* reader);
int memb_b200_device(const memb_b200_reader* reader);
memb_b200_status memb_b200_get_load_stats(const memb_b200_reader* reader, memb_b200_load_stats* stats);

/* Reader::keys() — src/reader.cpp:36-39, src/trained_compression.cpp:142-153.
 * words: NUL-terminated words back to back in row-id order; offsets has rows+1
 * entries.  Pointers stay valid until close(). */
memb_b200_status memb_b200_keys(const memb_b200_reader* reader, const char** words,
                                const uint32_t** offsets, uint64_t* rows);

/* The resolve half of CompressedStorage::extract — src/trained_compression.cpp:113-127,
 * flatbuffers LookupByKey in src/uniform_compression.cpp:56 / src/full_compression.cpp:39.
 * ids_host[i] = row id or -1.  Runs on the device (vocabulary hash). */
memb_b200_status memb_b200_resolve(memb_b200_reader* reader, const char* query_bytes,
                                   const uint64_t* query_offsets, uint64_t n, int32_t* ids_host);
/* The same with every buffer already on the reader's device (additive): query text, n+1
 * offsets and the n ids are DEVICE pointers; the resolve kernel is queued on `stream` and
 * the call returns without waiting (feeds memb_b200_decode_ids / union_decode_ids with
 * device ids; lets a caller time K0 alone). */
memb_b200_status memb_b200_resolve_device(memb_b200_reader* reader, const char* query_bytes_device,
                                          const uint64_t* query_offsets_device, uint64_t n,
                                          int32_t* ids_device, void* stream);

/* Reader::batchEmbeddingToBuffer(words, buffer) — src/reader.cpp:59-86.
 * out_host: n*dim floats, row-major. */
memb_b200_status memb_b200_lookup_words(memb_b200_reader* reader, const char* query_bytes,
                                        const uint64_t* query_offsets, uint64_t n, float* out_host);
/* Same lookup with the result left where the caller wants it (additive; SURVEY 8(f)-2:
 * python/memb/reader.py:94-111 feeds a GPU embedding layer).  out: host or device memory
 * on the reader's device; complete on return. */
memb_b200_status memb_b200_lookup_words_to(memb_b200_reader* reader, const char* query_bytes,
                                           const uint64_t* query_offsets, uint64_t n, float* out,
                                           memb_b200_mem_kind out_kind);
/* The same into DEVICE memory with the caller's CUDA stream: the query upload, the resolve
 * and the decode are queued on `stream` (0 = the legacy default stream), i.e. AFTER whatever
 * the caller already queued there -- the ordering a stream-ordered allocator (torch's caching
 * allocator) relies on when it hands out a block that earlier kernels of that stream still
 * use.  Complete on return (the reader's query scratch is reused by the next call). */
memb_b200_status memb_b200_lookup_words_on(memb_b200_reader* reader, const char* query_bytes,
                                           const uint64_t* query_offsets, uint64_t n, float* out_device,
                                           void* stream);

/* The decode half of extract for n row ids — src/trained_compression.cpp:129-136,
 * src/uniform_compression.cpp:58-73, src/full_compression.cpp:41-42, plus the
 * zero fill of src/reader.cpp:43-46.  With both pointers on the device the call is
 * asynchronous on `stream` (a cudaStream_t, may be NULL); otherwise it returns
 * when out is complete. */
memb_b200_status memb_b200_decode_ids(memb_b200_reader* reader, const int32_t* ids,
                                      memb_b200_mem_kind ids_kind, uint64_t n, float* out,
                                      memb_b200_mem_kind out_kind, void* stream);

/* Words -> rows WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same result as memb_b200_lookup_words / _on, but every
 * buffer is the caller's and already on the reader's device: query text, n + 1 uint64 offsets, and the
 * n x dim float output.  The one piece of scratch the lookup needs -- the n resolved row ids that
 * resolve_words_kernel hands to the decode kernel -- lives in a file-scope `static __device__` array of
 * MEMB_B200_NOALLOC_MAX_WORDS ids (4 MiB, held by every GPU that loads the library until exit)
 * instead of the reader's cudaMalloc'ed, growing scratch.  n > MEMB_B200_NOALLOC_MAX_WORDS is rejected
 * with MEMB_B200_ERR_ARGUMENT.  The call performs no cudaMalloc / cudaFree / stream-ordered allocation;
 * both kernels are queued on `stream` and the call returns without waiting.
 * The array is SHARED by every call in the process (all readers of a device): ids [0, n) are rewritten
 * completely by each call and nothing in it has to start at zero, but two calls that overlap on the GPU
 * (different streams, or different threads) race on it -- order them on one stream or with events. */
#define MEMB_B200_NOALLOC_MAX_WORDS (1u << 20)
memb_b200_status memb_b200_lookup_words_noalloc(memb_b200_reader* reader, const char* query_bytes_device,
                                                const uint64_t* query_offsets_device, uint64_t n,
                                                float* out_device, void* stream);

/* Full-vocabulary path (Reader.tokenizer_embedding / to_keyed_vectors,
 * python/memb/reader.py:19-30,94-111): rows [row_begin, row_end) in row-id
 * order; a rank of a multi-GPU job passes its own slice. */
memb_b200_status memb_b200_decode_range(memb_b200_reader* reader, uint64_t row_begin, uint64_t row_end,
                                        float* out, memb_b200_mem_kind out_kind, void* stream);

/* Fused decode -> all-gather for the sharded full-vocabulary decode ("each GPU stores its
 * slice straight into the owner's buffer").  out_multicast is an NVLS MULTICAST address of the
 * gathered matrix's slot for these rows (e.g. torch symmetric memory's multicast_ptr + offset): the
 * decode kernel's stores are multimem.st, which the NVSwitch replicates into every GPU of the group,
 * so the transfer overlaps the decode and no separate collective moves the rows.  Asynchronous on
 * `stream`; the caller synchronises the group (a barrier on the symmetric memory) before reading.
 * Trained readers with dim % 4 == 0 only (MEMB_B200_ERR_UNSUPPORTED otherwise). */
memb_b200_status memb_b200_decode_range_multicast(memb_b200_reader* reader, uint64_t row_begin, uint64_t row_end,
                                                  float* out_multicast, void* stream);

/* ReadersUnion.batch_embedding — python/memb/readers_union.py:90-96 with
 * np.mean(axis=0) / np.concatenate(axis=-1) (:16-18,30-32) fused into the decode.
 * All readers must live on the same device.  ids: num_readers*n, reader-major.
 * out: n * union_dim floats (union_dim = dim for average, sum of dims for concatenate). */
memb_b200_status memb_b200_union_lookup_words(memb_b200_reader* const* readers, uint32_t num_readers,
                                              memb_b200_union_mode mode, const char* query_bytes,
                                              const uint64_t* query_offsets, uint64_t n, float* out_host);
memb_b200_status memb_b200_union_lookup_words_to(memb_b200_reader* const* readers, uint32_t num_readers,
                                                 memb_b200_union_mode mode, const char* query_bytes,
                                                 const uint64_t* query_offsets, uint64_t n, float* out,
                                                 memb_b200_mem_kind out_kind);
/* device output on the caller's stream: see memb_b200_lookup_words_on */
memb_b200_status memb_b200_union_lookup_words_on(memb_b200_reader* const* readers, uint32_t num_readers,
                                                 memb_b200_union_mode mode, const char* query_bytes,
                                                 const uint64_t* query_offsets, uint64_t n, float* out_device,
                                                 void* stream);
memb_b200_status memb_b200_union_decode_ids(memb_b200_reader* const* readers, uint32_t num_readers,
                                            memb_b200_union_mode mode, const int32_t* ids,
                                            memb_b200_mem_kind ids_kind, uint64_t n, float* out,
                                            memb_b200_mem_kind out_kind, void* stream);

/* Pinned host memory for query/out buffers (PCIe-rate H2D/D2H). */
memb_b200_status memb_b200_host_alloc(size_t bytes, void** ptr);
void memb_b200_host_free(void* ptr);

/* Kernels launched by this library in this process so far (bench's gpu_launches). */
uint64_t memb_b200_launch_count(void);
/* Knobs for experiments and tests: "trained_lut_replicas" = 0 auto | 1 | 32,
 * "trained_warps" = 0 auto | 1..10, "trained_generic" = 0|1 (one-thread-per-row decode),
 * "trained_tma" = 0 | 1 (default: bulk-async staging for aligned range decodes and long records) | 2 everywhere,
 * "trained_rows_kernel" = 0 (default) | 1 unions of trained readers in one launch | 2 single lookups too,
 * "trained_rows_warps" = 0 auto | cap on that kernel's resident warps,
 * "trained_fixed_records" = 0 | 1 (default: rows as fixed-stride records when the longest row is within 25 % of the
 *   average), "trained_record_align" = 1 (default) | 2 | 4 (record stride a multiple of this many 16 B units),
 * "union_symbol_average" = 0|1 (average of trained readers through symbol words),
 * "union_pair_table" = 0 two gathers | 1 | 2 one table copy | 3 (default) one copy per shared-memory bank,
 * "union_dedup" = 0 every query of a union lookup is decoded | 1 (default) batches of >= 131 072 queries that repeat
 *   their id tuples (Zipf-distributed words) have the DISTINCT tuples decoded once and the rows copied out when a cost
 *   model says that is cheaper -- same bytes; the decision takes two 4-byte device -> host round trips on the call's
 *   stream, so it is skipped under stream capture | 2 always (tests),
 * "union_dedup_keys", "union_dedup_keep", "union_dedup_symbols" = 0|1 (its claim-kernel flavour, the L2 hint of its
 *   copy-out, symbol words instead of fp32 rows as its temporary),
 * "uniform_rows_kernel" = 0 element walk | 1 auto (row-group kernel unless levels is a power of two) | 2 always,
 * "host_pass_rows" = rows staged per pass for host outputs (0 = 4 GiB of fp32).
 * Applies to readers opened / calls made afterwards. */
memb_b200_status memb_b200_set_option(const char* name, int64_t value);

/* ---- write side (fixture generator; `_memb.Builder`, python/memb_bindings.cpp:12-34) ---- */
/* Builder(dim, storageName, bitsPerWeight) — src/builder.cpp:26-32 */
memb_b200_status memb_b200_builder_create(uint64_t dim, const char* storage_name, uint64_t bits_per_weight,
                                          memb_b200_builder** out);
void memb_b200_builder_destroy(memb_b200_builder* builder);
/* Builder::addWord — src/builder.cpp:34-48 */
memb_b200_status memb_b200_builder_add_word(memb_b200_builder* builder, const char* word, size_t word_len,
                                            const float* values, uint64_t count);
/* Bulk form: n words as byte ranges of one blob, vectors as a row-major (n, dim) matrix. */
memb_b200_status memb_b200_builder_add_words(memb_b200_builder* builder, const char* blob,
                                             const uint64_t* offsets, uint64_t n, const float* matrix);
/* GPU converter (additive): run the bulk passes of the trained compressor --
 * predict every value (src/kmeans.cpp:65-79), Huffman-encode every row byte-aligned
 * (src/trained_compression.cpp:54-71) -- on CUDA device `device` at save(); -1 = host threads (default).
 * The k-means fit on the 10 000-row sample stays on the host (its running mean is order dependent,
 * src/kmeans.cpp:86-102).  The file is byte-identical either way. */
memb_b200_status memb_b200_builder_set_device(memb_b200_builder* builder, int device);
/* Builder::save — src/builder.cpp:64-68 */
memb_b200_status memb_b200_builder_save(memb_b200_builder* builder, const char* path, unsigned num_threads);
/* The Builder's 1-D k-means on its own (additive; host only, no device needed): KMeansClusterizer(clusters)
 * .fit(data) then .predict(data) — src/kmeans.cpp:26-79 — so that the reference's own k-means test
 * (src/kmeans_tests.cpp:9-38) can be replayed against it.  ids: n cluster indices; centroids: room for
 * `clusters` floats, *num_centroids = how many survived the pruning of small clusters. */
memb_b200_status memb_b200_kmeans_fit_predict(const float* data, uint64_t n, uint32_t clusters, uint8_t* ids,
                                              float* centroids, uint32_t* num_centroids);
/* availableCompressionStrategies() — src/compression_strategy.cpp:63-78; returns "full\0uniform\0trained\0\0" */
const char* memb_b200_available_compression_strategies(void);

#ifdef __cplusplus
}
#endif
#endif /* MEMB_B200_H */
