/*
 * tfp_store.h -- C ABI of libtfpstore.so, the B200 (sm_100a) implementation of
 * LocalAI's `local-store` vector store.
 *
 * This is the drop-in boundary: a Go host (go/local-store/store_gpu.go) binds
 * these entry points through cgo and keeps the reference's gRPC surface
 * (backend/backend.proto:24-27) unchanged.  Every entry point names the
 * reference interface it replaces (paths relative to the reference tree).
 *
 * Conventions (the reference's own cgo shims use the same ones,
 * backend/go/bark/gobark.h:1-7): every function returns int, 0 = ok;
 *   > 0  a validation error the reference would also report -- the text from
 *        tfp_last_error() is byte-identical to store.go's fmt.Errorf text;
 *   < 0  a CUDA / driver failure (text from tfp_last_error()).
 * There is no CPU fallback anywhere behind this header.
 *
 * Keys are fp32, row-major, `dim` floats per key.  Ids are int64 handles the
 * library assigns at insertion; the host keeps id -> value bytes (values never
 * cross this boundary, store.go:23).
 *
 * Threading: calls on one store may come from any thread; mutations are
 * serialised internally, finds on different find-contexts may overlap.
 */
#ifndef TFP_STORE_H
#define TFP_STORE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tfp_store tfp_store;
typedef struct tfp_find_ctx tfp_find_ctx;

typedef struct {
    int      device;         /* CUDA device ordinal this store (shard) lives on            */
    int64_t  capacity_rows;  /* initial HBM reservation in rows; 0 = default, grows on use */
    uint32_t flags;          /* TFP_FLAG_*                                                  */
    uint32_t reserved;
    int64_t  id_base;        /* first id this store hands out (shards of one logical store
                                use disjoint ranges, e.g. rank << 40)                      */
} tfp_cfg;

#define TFP_FLAG_NONE 0u

/* Largest top-k served by the fused in-kernel selection; larger k takes the
 * score-spill + radix-sort path (same results, more HBM traffic). */
#define TFP_FUSED_TOPK_MAX 1024

/* Which arithmetic a find used (store.go:496-510 dispatch). */
#define TFP_PATH_NORMALIZED 1   /* store.go:327-339, fp32 sequential dot          */
#define TFP_PATH_FALLBACK   0   /* store.go:414-428, fp64-accumulated cosine      */

/* ---- lifecycle -------------------------------------------------------------
 * replaces NewStore() (backend/go/stores/store.go:39-46) and Load() (:101-103). */
int  tfp_store_open(const tfp_cfg* cfg, tfp_store** out);
void tfp_store_close(tfp_store* s);

/* ONE logical store over several GPUs of one box, inside one process.
 * LocalAI spawns a single `local-store` process per backend and passes it only --addr and
 * the environment (pkg/model/initializers.go:333-368, pkg/model/process.go:94-124), so
 * the multi-GPU store sits behind the SAME handle and the SAME entry points as the
 * one-GPU store: every tfp_store_* function below accepts the handle this returns.
 * devices: n_devices CUDA ordinals (1..8; one shard per entry; cfg->device is ignored;
 * cfg->capacity_rows is the whole store's hint).  Keys are routed to a shard by a hash
 * of their canonical bits (tfp_route_rows); a find scans every shard, the scan kernels
 * write their result records into GPU devices[0]'s memory over NVLink (peer access;
 * without it the per-shard answers are merged on the host), one merge kernel there
 * produces the answer.  Calls may come from any thread; up to TFP_MULTI_LANES (default 4)
 * finds run at once, further callers wait their turn.  Shard g hands out ids from
 * cfg->id_base + (g << 40).  tfp_store_set_device works on such a handle but is not
 * zero-copy there (rows must be routed: the block travels through page-locked host
 * memory).  Not available: the find-context calls (tfp_find_ctx_create fails; the store
 * runs its own lanes). */
int  tfp_store_open_devices(const tfp_cfg* cfg, const int* devices, int n_devices, tfp_store** out);
/* Number of shards behind a handle (1 for tfp_store_open) and one shard's figures: what a
 * host's Status reply lists per GPU (pkg/grpc/base/base.go:77-81 memory breakdown). */
int  tfp_store_shard_count(tfp_store* s);
/* Which shard of an n_shards-wide store owns each of n keys (pure host arithmetic; equal
 * keys under Go's cmp.Compare -- -0 == +0, NaN == NaN, store.go:48-52 -- share a shard). */
int  tfp_route_rows(const float* keys, int64_t n, int dim, int n_shards, int* out_shard);

/* Text of the calling thread's most recent non-zero return.  Thread-local: a host whose
 * calls may change OS thread between the failing call and this one (goroutines under
 * cgo) either pins the thread around the pair (runtime.LockOSThread) or uses
 * tfp_store_last_error. */
const char* tfp_last_error(void);
/* Text of the most recent failing call ON THIS HANDLE, whichever thread made it, copied
 * into buf (NUL-terminated, truncated to cap).  SURVEY App. D tfp_store_last_error. */
int tfp_store_last_error(tfp_store* s, char* buf, size_t cap);

/* ---- StoresSet: replaces Store.StoresSet (store.go:106-198) ------------------
 * keys: n x dim floats in host memory, copied before return (the Go host may
 * not lend Go pointers past the call, cgo rules).  n_values is len(opts.Values)
 * and is only checked against n (store.go:111-113).
 * out_ids[i]: id now holding key i; out_was_new[i]: 1 = appended, 0 = an equal
 * stored key was replaced (store.go:183-187; equality is Go's cmp.Compare:
 * -0 == +0, NaN == NaN).  Duplicates inside one call are both stored, as in the
 * reference (each stored key is replaced at most once per call). */
int tfp_store_set(tfp_store* s, const float* keys, int64_t n, int dim, int64_t n_values,
                  int64_t* out_ids, uint8_t* out_was_new);

/* Same, with the keys already in this device's HBM (bulk load / ingest from an
 * embedding model's output).  No reference counterpart below the API: the
 * reference has no device.  Same semantics as tfp_store_set. */
int tfp_store_set_device(tfp_store* s, const float* d_keys, int64_t n, int dim,
                         int64_t* out_ids, uint8_t* out_was_new);

/* ---- StoresDelete: replaces Store.StoresDelete (store.go:200-264) ------------
 * out_ids[i] = id removed for key i, or -1 when the key was not stored
 * (missing keys are ignored, as in the reference). */
int tfp_store_delete(tfp_store* s, const float* keys, int64_t n, int dim, int64_t* out_ids);

/* ---- StoresGet: replaces Store.StoresGet (store.go:266-311) ------------------
 * out_ids[i] = id for request key i or -1.  out_order[0..*out_n_found) lists
 * the request indices of the found keys in sorted-key order, the order the
 * reference returns them in (store.go:269,285-301).
 * Deviation, documented: n == 0 returns an error; the reference panics
 * (store.go:276 indexes opts.Keys[0]). */
int tfp_store_get(tfp_store* s, const float* keys, int64_t n, int dim,
                  int64_t* out_ids, int64_t* out_order, int64_t* out_n_found);

/* ---- StoresFind: replaces Store.StoresFind / StoresFindNormalized /
 * StoresFindFallback (store.go:477-511, 373-412, 430-475) -----------------------
 * q: dim floats, host memory.  Results in descending similarity:
 * out_ids/out_sims hold *out_n = min(topk, rows) entries (caller provides topk
 * slots); out_keys, when not NULL, receives the matching key rows (topk x dim),
 * which StoresFindResult.Keys needs.  out_path, when not NULL, gets TFP_PATH_*.
 * Scores are the reference's arithmetic bit for bit (amd64, no FMA); among
 * exactly equal scores the lower slot (older row) ranks first. */
int tfp_store_find(tfp_store* s, const float* q, int dim, int topk,
                   int* out_n, int64_t* out_ids, float* out_sims, float* out_keys, int* out_path);

/* nq independent finds that share passes over the keys.  Same per-query results
 * as nq calls of tfp_store_find (bit-identical similarities).  Up to 4 same-path
 * queries share one pass of the exact scan kernel; 3 or more on a store of
 * >= 65536 rows (topk <= 100) go through the tensor cores in groups of up to 256
 * (TF32 filter + exact re-score, mma_scan.cuh).  Replaces nq sequential
 * Store.StoresFind calls (the reference has no batch RPC).
 * out_n[nq]; out_ids/out_sims [nq x topk]; out_keys [nq x topk x dim] or NULL. */
int tfp_store_find_batch(tfp_store* s, const float* q, int nq, int dim, int topk,
                         int* out_n, int64_t* out_ids, float* out_sims, float* out_keys);

/* ---- introspection: feeds Status / memory breakdown (pkg/grpc/base/base.go:77-81) */
typedef struct {
    int64_t rows;
    int64_t capacity_rows;
    int64_t hbm_bytes;
    int     dim;                  /* -1 until latched (store.go:29,44)              */
    int     keys_are_normalized;  /* sticky flag, store.go:27,43,126-127            */
    int     device;
    int     reserved;
} tfp_stats;
int tfp_store_stats(tfp_store* s, tfp_stats* out);
int tfp_store_shard_stats(tfp_store* s, int shard, tfp_stats* out);

/* ---- snapshot / restore ----------------------------------------------
 * The reference keeps the store in process memory only (Load is a no-op, store.go:101-103): a
 * watchdog kill or a respawn loses everything.  tfp_store_save writes rows, ids, key length, the
 * sticky flag and the id counter to `path`; tfp_store_load fills an EMPTY store from such a file
 * (hashes and norms are recomputed on the device).  Values stay with the host, keyed by id. */
int tfp_store_save(tfp_store* s, const char* path);
int tfp_store_load(tfp_store* s, const char* path);
/* Crash consistency.  A save writes `path.tmp`, fsyncs it and renames it over `path`: a kill at any
 * moment leaves either the previous snapshot or the complete new one (a multi-GPU store writes one
 * file per shard, `path.g<i>`, and its manifest at `path` last; a load checks the shards against it).
 * The host's values live in a file of the host's own; to tie the two together the host picks a fresh
 * 64-bit tag per save, writes its values to a NEW file named after the tag, then calls
 * tfp_store_save_tagged -- the rename of `path` is the commit point -- and only then removes older
 * values files.  At start it reads the tag back with tfp_snapshot_tag and opens the values file of
 * that name.  tfp_store_save is tfp_store_save_tagged with tag 0. */
int tfp_store_save_tagged(tfp_store* s, const char* path, uint64_t tag);
int tfp_snapshot_tag(const char* path, uint64_t* out_tag);

/* A shard of a logical multi-GPU store is told that another shard has seen a key that
 * is not unit length: the sticky flag (store.go:27,126-127) is a property of the whole
 * store, so every shard must leave the normalised path together. */
int tfp_store_mark_not_normalized(tfp_store* s);

/* ---- find contexts: the asynchronous, device-resident face of StoresFind ------
 * A context owns a CUDA stream, staged queries in HBM, scratch for the
 * per-CTA candidate lists and device-resident results.  It exists so that
 * (a) several finds can be in flight, (b) a multi-GPU host can all-gather the
 * per-shard candidates without a host round trip, and (c) the
 * scan can be timed with inputs already in HBM.  No reference counterpart:
 * the reference is synchronous and single-threaded (store.go:18).        */
int  tfp_find_ctx_create(tfp_store* s, tfp_find_ctx** out);
/* The context behind tfp_store_find / tfp_store_find_batch (owned by the store; do not destroy):
 * lets a host read its counters. */
tfp_find_ctx* tfp_store_default_ctx(tfp_store* s);
void tfp_find_ctx_destroy(tfp_find_ctx* c);
/* Stage nq host queries into HBM (async on the context's stream); computes each
 * query's path and |q| on the host exactly as store.go:313-324,435-439. */
int  tfp_find_ctx_upload(tfp_find_ctx* c, const float* q, int nq, int dim);
/* Enqueue the scan + top-k for staged queries [first, first+count).  Results
 * stay in HBM.  Returns without waiting for the GPU.  A batch that takes the
 * tensor-core path (>= 3 same-path queries on a large store) is finished in two
 * steps: its per-query status words decide whether some queries need a second
 * pass (3xTF32 filter, then the exact kernel), and somebody has to read them.
 * When the launch is ONE such batch (count <= 256, one path) this call enqueues
 * the first pass and returns; the NEXT call on the context -- upload, launch,
 * gather_keys, download, sync, device_results, device_records -- waits for the
 * status words and enqueues what is left, before it does its own work.  So the
 * answer is final after any of those calls has returned and the stream has
 * drained; a host that orders other work after the context's stream must call
 * tfp_find_ctx_sync (not only wait for the stream) first.  A StoresSet /
 * StoresDelete that has to move rows finishes such batches itself, on the rows
 * as they were.  Launches made of several batches, and the synchronous and
 * sharded entry points, read the status words before they return.
 * (TFP_MMA_DEFER=0: always read them before returning.) */
int  tfp_find_ctx_launch(tfp_find_ctx* c, int first, int count, int topk);
/* Wait, then copy results of the last launch to the host (same layout as
 * tfp_store_find_batch for `count` queries). */
int  tfp_find_ctx_download(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims, float* out_keys);
/* Wait for everything the last launch needs, later tensor-core passes included. */
int  tfp_find_ctx_sync(tfp_find_ctx* c);
/* The context's cudaStream_t, for event timing and for ordering collectives. */
void* tfp_find_ctx_stream(tfp_find_ctx* c);
/* Device pointers of the last launch's results: cand[count x topk] packed
 * (score-order word << 32 | ~slot), ids[count x topk], sims[count x topk], n[count].
 * Returns the entries per query in those arrays (min(topk, rows)), negative on failure. */
int  tfp_find_ctx_device_results(tfp_find_ctx* c, uint64_t** d_cand, int64_t** d_ids,
                                 float** d_sims, int** d_n);
/* Enqueue the gather of the last launch's winning key rows and return their device
 * address: [count x stride x dim] floats, stride as returned by
 * tfp_find_ctx_device_results (what StoresFindResult.Keys is built from). */
int  tfp_find_ctx_gather_keys(tfp_find_ctx* c, float** d_keys);
/* How often this context ran a tensor-core pass over the keys (batches of >= 3 same-path queries on a
 * store of >= 65536 rows: tcgen05 single-TF32 filter on CTA pairs + exact re-score) and how many queries
 * it then had to re-run on the exact scan kernel.  Either pointer may be NULL.
 * tfp_find_ctx_mma_escalated: queries whose scores lay too close together for the single-TF32 filter
 * and went through a second pass with the 3xTF32 filter (counted in `passes` too). */
int  tfp_find_ctx_mma_counters(tfp_find_ctx* c, int64_t* passes, int64_t* exact_reruns);
int64_t tfp_find_ctx_mma_escalated(tfp_find_ctx* c);
/* Number of kernels the context has launched so far (for bench accounting). */
int64_t tfp_find_ctx_launch_count(tfp_find_ctx* c);

/* ---- cross-shard merge (no reference counterpart) ------------------
 * Every launch also writes one fixed-size RECORD per query, so that a multi-GPU host
 * can move a shard's whole answer with ONE all-gather:
 *     int32 n; int32 reserved; int64 ids[topk]; float sims[topk]; (padded to 16 bytes)
 * topk is the caller's topk (not min(topk, rows)), so records have the same size on
 * every shard.  tfp_record_bytes gives that size. */
size_t tfp_record_bytes(int topk);
/* Device address of the last launch's records ([count] records) and their size. */
int tfp_find_ctx_device_records(tfp_find_ctx* c, void** d_records, size_t* record_bytes);
/* After an all-gather of every shard's records for one launch (layout
 * [n_shards x nq] records in HBM), pick the global top-k per query on the device.
 * Outputs [nq x topk] / [nq] in HBM; also reports the winning shard and the index
 * inside that shard's list.  Ties: lower shard, then that shard's own order.
 * stream is a cudaStream_t (NULL = default stream). */
int tfp_merge_records(const void* d_gathered, int n_shards, int nq, int topk,
                      float* d_out_sims, int64_t* d_out_ids, int* d_out_shard,
                      int* d_out_index, int* d_out_n, void* stream);

/* ---- sharded find in one call (no reference counterpart) --------------
 * The exchange step of a multi-GPU find -- one ncclAllGather of every shard's fixed-size
 * record over NVLink, then the device merge -- enqueued by the library itself on the
 * context's stream, so that a host pays one call per query instead of three.
 * NCCL is resolved at run time from the process (dlopen of libnccl.so.2: the copy the
 * host already uses, e.g. the one bundled with PyTorch, else the system one); the library
 * does not link against it.  One communicator per find context: contexts are independent
 * lanes, and NCCL serialises what shares a communicator.
 *   tfp_nccl_unique_id   rank 0 creates the 128-byte id and hands it to the other ranks
 *                        by whatever means the host has (torch.distributed, MPI, a file).
 *   tfp_find_ctx_comm_init   collective: every rank calls it with the same id.
 *   tfp_find_ctx_launch_sharded  scan this shard for staged query `qi`, all-gather the
 *                        records, merge.  Returns without waiting.  Every rank must issue
 *                        the same sequence of sharded launches per context.
 *   tfp_find_ctx_merged_results  device pointers of the merged answer ([count x topk] each, n[count]).
 *   tfp_find_ctx_download_merged wait and copy the merged answer to the host (stride topk per query). */
int tfp_nccl_unique_id(void* out_id, size_t out_bytes);
int tfp_find_ctx_comm_init(tfp_find_ctx* c, const void* unique_id, int rank, int world);
/* Let `c` use the communicator `owner` created (both contexts of one store).  NCCL then orders
 * their collectives one after the other; `owner` must outlive `c`'s last sharded launch. */
int tfp_find_ctx_comm_share(tfp_find_ctx* c, tfp_find_ctx* owner);
int tfp_find_ctx_launch_sharded(tfp_find_ctx* c, int qi, int topk);
/* Same for staged queries [first, first+count): one all-gather of count records, one merge CTA per
 * query.  Batches of >= 3 same-path queries take the tensor-core path on every shard. */
int tfp_find_ctx_launch_sharded_batch(tfp_find_ctx* c, int first, int count, int topk);
int tfp_find_ctx_merged_results(tfp_find_ctx* c, float** d_sims, int64_t** d_ids, int** d_shard,
                                int** d_index, int** d_n);
int tfp_find_ctx_download_merged(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims,
                                 int* out_shard, int* out_index);

/* ---- sharded find over peer memory (no reference counterpart) ----------
 * The same exchange without a collective library call: the scan kernel's last CTA writes this
 * shard's records straight into EVERY rank's gather block over NVLink (peer-mapped memory) and
 * raises a sequence word there; the receiving context's stream waits for its peers' words with a
 * stream memory operation (cuStreamWaitValue32 -- no kernel spins) and merges.  One process per
 * GPU, 2..8 GPUs with peer access (one NVSwitch box).  Set-up, per find context:
 *   tfp_find_ctx_p2p_export  allocate this rank's gather block (2 parities x world slots of
 *                        slot_bytes, + world sequence words) and return its 64-byte cudaIpc handle;
 *   (the host hands every rank's handle to every rank: torch.distributed, MPI, a file)
 *   tfp_find_ctx_p2p_attach  map the peers' blocks; `handles` = world x 64 bytes in rank order.
 *                        Every rank must have attached before any rank launches (barrier).
 * After that tfp_find_ctx_launch_sharded[_batch] uses the peer exchange whenever count records fit
 * slot_bytes, else the NCCL communicator if one was set.  Every rank must issue the same sequence
 * of sharded launches per context (a rank that skips one leaves its peers' streams waiting).
 *   tfp_find_ctx_p2p_detach  unmap the peers' blocks.  Every rank detaches, THEN (barrier) contexts
 *                        are destroyed: an exported block must not be freed while a peer maps it. */
int tfp_find_ctx_p2p_export(tfp_find_ctx* c, int world, size_t slot_bytes, void* handle_out, size_t handle_bytes);
int tfp_find_ctx_p2p_attach(tfp_find_ctx* c, const void* handles, int rank, int world);
int tfp_find_ctx_p2p_detach(tfp_find_ctx* c);
/* The same find, with the winning key rows (StoresFindResult.Keys, store.go:399-401): every shard's
 * top-k rows travel to every rank with the records, the merged answer carries the rows of the global
 * winners, and tfp_find_ctx_download_merged_keys brings n, ids, sims and rows to the host in ONE copy
 * (stride topk per query; out_keys [count x topk x dim]).  Peer exchange only; the slot must hold
 * tfp_p2p_slot_bytes_with_keys(count, topk, dim) bytes. */
size_t tfp_p2p_slot_bytes_with_keys(int count, int topk, int dim);
int tfp_find_ctx_launch_sharded_keys(tfp_find_ctx* c, int first, int count, int topk);
int tfp_find_ctx_download_merged_keys(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims,
                                      float* out_keys);

/* ---- pinned host staging (no reference counterpart) --------------------------------
 * Keys handed to tfp_store_set / queries handed to the find entry points are copied to
 * the GPU from wherever they are; from ordinary (pageable) memory that copy runs at
 * ~7 GB/s.  A host that fills its request buffers directly in memory obtained here
 * (page-locked) gets the full PCIe rate.  Plain malloc/free semantics. */
void* tfp_host_alloc(size_t bytes);
void  tfp_host_free(void* p);

/* ---- a find that allocates nothing ----------------------------------------------
 * This entry point has no counterpart in the reference.
 *
 * Some hosts must not cause a device allocation while they serve requests: a
 * runtime that has already handed all of HBM to its own arena (an LLM or
 * embedding server sharing the GPU with the store) has nothing left for the
 * driver to give, and a latency-critical loop cannot afford cudaMalloc /
 * cudaFree, which synchronise the whole device.  Every other find in this
 * header may grow a buffer the first time it sees a larger top-k or query count.
 * This one never does: it calls no cudaMalloc, cudaFree or allocating library
 * routine.  It runs the same scan + fused top-k kernel as tfp_store_find on the
 * normalised path (fp32 index-order dot, store.go:327-339) over a key matrix
 * the CALLER owns, and its scratch (per-CTA candidate lists, counters, the
 * packed record) lives in a fixed array inside the library image:
 * TFP_NOALLOC_SCRATCH_BYTES of device memory that every GPU which loads the
 * library holds until the process exits.
 *
 * d_keys [n_rows x dim] fp32 row-major, d_ids [n_rows], d_query [dim]: inputs in
 * the current device's memory.  d_out_sims / d_out_ids [topk], d_out_n [1]:
 * outputs in device memory, written in stream order (descending similarity,
 * *d_out_n = min(topk, n_rows) valid entries, ties: lower row first).
 * Limits, checked and reported through tfp_last_error(): 1 <= topk <=
 * TFP_NOALLOC_TOPK_MAX, 1 <= dim <= TFP_NOALLOC_DIM_MAX, n_rows < 2^32 - 16.
 * The first call on each device looks up the scratch array's device address
 * and opts the kernels into large shared memory; make that first call with the
 * device current and before starting any stream capture.  All calls on one
 * device share the one scratch array: the library clears the words that must
 * start at zero on every call, but two calls whose kernels overlap on the GPU
 * (different streams, no ordering between them) race on it -- order them.
 * stream is a cudaStream_t (NULL = default stream); the call returns without
 * waiting for the GPU. */
#define TFP_NOALLOC_TOPK_MAX 64
#define TFP_NOALLOC_DIM_MAX 8192
#define TFP_NOALLOC_SCRATCH_BYTES (672 * 1024)
int tfp_scan_topk_noalloc(const float* d_keys, int64_t n_rows, int dim, const int64_t* d_ids,
                          const float* d_query, int topk,
                          float* d_out_sims, int64_t* d_out_ids, int* d_out_n, void* stream);

/* Build/version string (contains the SM target the kernels were built for). */
const char* tfp_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TFP_STORE_H */
