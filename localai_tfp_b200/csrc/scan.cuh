// scan.cuh -- the StoresFind hot loop on sm_100a.
//
// Replaces `for i, k := range s.keys { sim := ...; heap.Push ...}` of
// StoresFindNormalized / StoresFindFallback
// (backend/go/stores/store.go:378-389, :441-452) and the two similarity
// functions (store.go:327-339 fp32 sequential dot; :414-428 fp64-accumulated
// cosine of fp32-rounded products).
//
// Design:
//   * persistent grid, one CTA per SM; tile = 128 consecutive rows;
//   * a producer warp streams each tile as [128 rows x 32 floats] TMA boxes
//     (128-byte swizzle) through a ring of shared-memory stages guarded by
//     full/empty mbarriers -- the key matrix is read from HBM exactly once;
//   * 128 consumer threads, ONE THREAD PER ROW: each thread walks its row in
//     index order with separately rounded multiply and add, i.e. exactly the
//     reference's arithmetic, so scores are bit-identical to the Go code.  The
//     swizzle makes the per-thread 16-byte shared loads bank-conflict free;
//   * NQ (1..8) queries share one pass over the keys;
//   * the top-k selection is fused (select.cuh); per-CTA survivors go to a small
//     scratch area in HBM and the last CTA to finish merges them.
#pragma once
#include "select.cuh"

namespace tfp {

constexpr int kBoxCols = 32;                               // floats per row per TMA box (128 B)
constexpr int kStageBytes = kRowsPerTile * kBoxCols * 4;   // 16 KiB
constexpr int kScanThreads = kRowsPerTile + 64;            // consumers + producer warp + bound-refresher warp
constexpr int kMaxStages = 12;
constexpr int kMaxNQ = 8;
constexpr int kMaxGrid = 512;                              // upper bound on CTAs per scan launch

struct QInfo {        // per staged query, filled on the host (exact reference arithmetic)
    double mag1;      // store.go:435-439
    int normalized;   // store.go:313-324 && store flag
    int pad;
};

// Peer exchange of result records over NVLink: the CTA that finishes a find writes
// this shard's packed records straight into every rank's gather buffer (peer-mapped memory) and
// then raises a sequence word there; the receiving rank's stream waits on that word (a stream
// memory operation, no kernel spins) before its merge kernel runs.  world == 0: no exchange.
constexpr int kXchgMaxWorld = 8;
struct XchgParams {
    uint8_t* slot[kXchgMaxWorld];        // where THIS rank's records land on rank r (r == me: the local buffer)
    unsigned int* flag[kXchgMaxWorld];   // rank r's "records from rank `me` have landed" word
    int world, me;
    unsigned int seq;                    // value to raise the words to (monotonic per find context)
    int raise;                           // this launch completes the find: raise the words after the copy
};

struct ScanParams {
    const float* keys;        // [capacity x dim] row-major
    const int64_t* ids;       // [capacity]
    const double* knorm;      // [capacity] sqrt(sum float64(float32(k*k))), store.go:420,423
    uint32_t n_rows;
    int dim;
    int n_chunks;             // ceil(dim / 32)
    int n_stages;
    int topk;                 // <= TFP_FUSED_TOPK_MAX on the fused path
    int cap;                  // selection buffer entries per query (power of two)
    int evict_first;          // stream keys through L2 with evict-first
    unsigned long long* dbg_times;   // development only: [grid x 4] globaltimer stamps (start, scan end, epilogue end, tiles), or nullptr
    const float* q;           // [nq_total x dim] staged queries
    const QInfo* qinfo;       // [nq_total]
    int q_first;              // first staged query of this launch group
    int out_first;            // first output slot of this launch group
    // scratch (per find-context)
    uint64_t* part;           // [grid x kMaxNQ x topk] per-CTA survivors
    int* part_cnt;            // [grid x kMaxNQ]
    unsigned long long* gtau; // [kMaxNQ] grid-wide admission bounds
    unsigned int* ticket;     // [1]
    unsigned int* tile_counter; // [1] dynamic tile scheduler: next tile = gridDim.x + counter++
    unsigned long long* lbest; // [kMaxNQ x lbest_stride] per-CTA "m-th best so far" (0 = not published)
    int lbest_stride;
    int m_local;              // m = ceil(topk / grid): which of its bests every CTA publishes
    int r_local;              // r = ceil(topk / m): the r-th largest published value is the bound
    int refresh_ns;           // refresher warp's poll period; 0 = no refresher
    uint64_t* all_cand;       // big-k path: [kMaxNQ? no: 1 x n_rows] every row's candidate word
    // outputs, indexed by out_first + query
    uint64_t* out_cand;       // [* x topk]
    int64_t* out_ids;         // [* x topk]
    float* out_sims;          // [* x topk]
    uint32_t* out_slots;      // [* x topk]
    int* out_n;               // [*]
    uint8_t* out_rec;         // [*] packed records (include/tfp_store.h), rec_bytes each, rec_topk slots
    int rec_bytes;
    int rec_topk;
    int qsel[8];              // use_qsel: the staged queries this launch serves (any subset, e.g. the queries the tensor-core
    int use_qsel;             //   filter could not settle); output slot of query j = out_first + (qsel[j] - q_first)
    XchgParams xchg;          // fused push of the finished records to the peer ranks (world == 0: off)
    float* out_keys;          // [* x topk x dim] the winning key rows (StoresFindResult.Keys), or nullptr: gathered by a later kernel
};

// Copy `bytes` (a multiple of 16) of finished records from local memory to the same offset of this
// rank's slot on every rank, then make the copies visible system-wide.  All `nthreads` threads call it;
// the caller puts a CTA barrier between this and xchg_raise.
__device__ __forceinline__ void xchg_copy(const XchgParams& x, const uint8_t* src, size_t off, size_t bytes, int tid,
                                          int nthreads) {
    const size_t words = bytes >> 4;
    const uint4* s = (const uint4*)src;
    for (size_t i = (size_t)tid; i < words * (size_t)x.world; i += (size_t)nthreads) {
        const int r = (int)(i / words);
        const size_t w = i - (size_t)r * words;
        const uint4 v = __ldcg(s + w);
        uint4* d = (uint4*)(x.slot[r] + off) + w;
        asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(d), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
    __threadfence_system();
}
__device__ __forceinline__ void xchg_raise(const XchgParams& x, int tid) {
    if (x.raise && tid < x.world && tid != x.me)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(x.flag[tid]), "r"(x.seq) : "memory");
}
// Stand-alone push for the paths whose records are written by other kernels (tensor-core finalize,
// big-k decode, multi-launch batches, empty shards).  One CTA per destination rank: it copies the
// records there and raises that rank's word itself, so no CTA waits for another.
__global__ void push_records_kernel(const __grid_constant__ XchgParams x, const uint8_t* src, size_t bytes) {
    const int r = (int)blockIdx.x;
    const size_t words = bytes >> 4;
    const uint4* s = (const uint4*)src;
    uint4* d = (uint4*)x.slot[r];
    for (size_t w = threadIdx.x; w < words; w += blockDim.x) {
        const uint4 v = __ldcg(s + w);
        asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(d + w), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0 && x.raise && r != x.me)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(x.flag[r]), "r"(x.seq) : "memory");
}

// Winning key rows of a sharded find, from the store straight into every rank's slot: CTA (i, q) copies
// row out_slots[q][i] to byte offset keys_off + (q * topk + i) * dim * 4 of this rank's slot everywhere.
// The words are raised by the push_records_kernel that follows in the stream.
__global__ void push_rows_kernel(const __grid_constant__ XchgParams x, const float* __restrict__ keys, int dim,
                                 const uint32_t* __restrict__ out_slots, const int* __restrict__ out_n, int stride, int topk,
                                 size_t keys_off) {
    const int i = blockIdx.x, q = blockIdx.y;
    if (i >= out_n[q]) return;
    const float* src = keys + (size_t)out_slots[(size_t)q * stride + i] * dim;
    const size_t off = keys_off + ((size_t)q * topk + i) * (size_t)dim * 4;
    for (int t = threadIdx.x; t < dim * x.world; t += blockDim.x) {
        const int r = t / dim, cidx = t - r * dim;
        ((float*)(x.slot[r] + off))[cidx] = src[cidx];
    }
    __threadfence_system();
}
// After the merge: winner i of query q sits in shard[q][i]'s slot at position index[q][i].
__global__ void gather_merged_rows_kernel(const uint8_t* block, size_t shard_stride, size_t keys_off, int dim, int topk,
                                          const int* __restrict__ shard, const int* __restrict__ index,
                                          const int* __restrict__ n, float* __restrict__ out) {
    const int i = blockIdx.x, q = blockIdx.y;
    if (i >= n[q]) return;
    const size_t at = (size_t)q * topk + i;
    const float* src = (const float*)(block + (size_t)shard[at] * shard_stride + keys_off) + ((size_t)q * topk + index[at]) * dim;
    float* dst = out + at * dim;
    for (int t = threadIdx.x; t < dim; t += blockDim.x) dst[t] = __ldcg(src + t);
}

// Staged query served as query j of a launch, and the output slot its answer goes to.
__device__ __forceinline__ int scan_query_of(const ScanParams& p, int j) { return p.use_qsel ? p.qsel[j] : p.q_first + j; }
__device__ __forceinline__ int scan_slot_of(const ScanParams& p, int j) { return p.out_first + (scan_query_of(p, j) - p.q_first); }

// ---- per-row arithmetic ----------------------------------------------------
template <bool F64> struct Acc;
template <> struct Acc<false> {
    float v;
    __device__ __forceinline__ void zero() { v = 0.0f; }
    // dot += k1[i] * k2[i]   (store.go:332) -- product and sum rounded separately
    __device__ __forceinline__ void add(float q, float k) { v = __fadd_rn(v, __fmul_rn(q, k)); }
};
template <> struct Acc<true> {
    double v;
    __device__ __forceinline__ void zero() { v = 0.0; }
    // dot += float64(k1[i] * k2[i])   (store.go:419)
    __device__ __forceinline__ void add(float q, float k) { v = __dadd_rn(v, (double)__fmul_rn(q, k)); }
};

template <bool F64>
__device__ __forceinline__ float finish_score(const Acc<F64>& a, double mag1, double knorm) {
    if constexpr (F64) {
        // sim := float32(dot / (mag1 * math.Sqrt(mag2)))   (store.go:423)
        return __double2float_rn(__ddiv_rn(a.v, __dmul_rn(mag1, knorm)));
    } else {
        return a.v;
    }
}

// ---- shared-memory carve-up --------------------------------------------------
struct ScanSmem {
    uint8_t* stages;             // n_stages x kStageBytes, 1024-aligned
    float* q;                    // NQ x dpad
    uint64_t* sel;               // NQ x cap
    uint64_t* full;              // n_stages
    uint64_t* empty;             // n_stages
    uint64_t* qbar;              // 1
    unsigned long long* tau;     // NQ
    int* cnt;                    // NQ
    int* flag;                   // 1
    int* pref;                   // kMaxGrid + 1 (last-CTA merge: prefix of survivor counts)
    uint32_t* tile_id;           // n_stages: which tile the stage's first chunk belongs to
    unsigned long long* ext;     // NQ: grid-wide bound kept fresh by the refresher warp
    unsigned long long* lmax;    // NQ: best word this CTA has seen (published as lbest when m == 1)
    int* done;                   // 1: consumers are finished, refresher may exit
    unsigned long long* lstage;  // kMaxGrid: the refresher warp's copy of the per-CTA bests
};
// Packed per-query record: int32 n; int32 reserved; int64 ids[topk]; float sims[topk]; padded to 16.
__host__ __device__ inline size_t record_bytes(int topk) {
    return ((size_t)8 + (size_t)topk * 12 + 15) & ~(size_t)15;
}
__device__ __forceinline__ void record_write(uint8_t* rec, int rec_topk, int i, int64_t id, float sim) {
    ((int64_t*)(rec + 8))[i] = id;
    ((float*)(rec + 8 + (size_t)rec_topk * 8))[i] = sim;
}

__host__ __device__ inline size_t scan_smem_bytes(int n_stages, int nq, int dpad, int cap) {
    size_t b = (size_t)n_stages * kStageBytes;
    b += (size_t)nq * dpad * 4;
    b = (b + 15) & ~(size_t)15;
    b += (size_t)nq * cap * 8;
    b += (size_t)(2 * n_stages + 1) * 8;
    b += (size_t)nq * 8 + (size_t)nq * 4 + 16 + (size_t)(kMaxGrid + 1) * 4 + (size_t)kMaxStages * 4;
    b += (size_t)nq * 16 + 16 + (size_t)kMaxGrid * 8 + 8;
    return b + 1024;  // slack for the 1024-byte alignment of the stage ring
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ ScanSmem scan_carve(uint8_t* raw, int n_stages, int nq, int dpad, int cap) {
    ScanSmem s;
    // pointer arithmetic only (no integer round trip), so the compiler keeps the
    // shared address space and emits LDS/STS rather than generic LD/ST
    uint8_t* p = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
    s.stages = p; p += (size_t)n_stages * kStageBytes;
    s.q = (float*)p; p += ((size_t)nq * dpad * 4 + 15) & ~(size_t)15;
    s.sel = (uint64_t*)p; p += (size_t)nq * cap * 8;
    s.ext = (unsigned long long*)p; p += (size_t)nq * 8;
    s.lmax = (unsigned long long*)p; p += (size_t)nq * 8;
    s.full = (uint64_t*)p; p += (size_t)n_stages * 8;
    s.empty = (uint64_t*)p; p += (size_t)n_stages * 8;
    s.qbar = (uint64_t*)p; p += 8;
    s.tau = (unsigned long long*)p; p += (size_t)nq * 8;
    s.cnt = (int*)p; p += (size_t)nq * 4;
    s.flag = (int*)p; p += 4;
    s.pref = (int*)p; p += (size_t)(kMaxGrid + 1) * 4;
    s.tile_id = (uint32_t*)p; p += (size_t)kMaxStages * 4;
    s.done = (int*)p; p += 4;
    p += (8u - (smem_u32(p) & 7u)) & 7u;
    s.lstage = (unsigned long long*)p;
    return s;
}

// The grid-wide bound from the per-CTA "m-th best so far" words of query q, by one warp: the r-th
// largest of them (r = ceil(k / m)).  r CTAs each hold m rows at least that good, so r*m >= k rows
// of the store reach it: it is a lower bound of the global k-th best.  For k <= #CTAs this is simply
// the k-th largest of the CTAs' best rows -- within a few ranks of the true k-th best, so almost
// nothing but the answer survives to the final merge.  `stage`: shared scratch for gridDim.x words.
// Returns 0 while fewer than r CTAs have published (then there is no bound yet).
__device__ __forceinline__ unsigned long long lbest_rth(const ScanParams& p, int q, int lane,
                                                        unsigned long long* stage) {
    const int G = (int)gridDim.x;
    const volatile unsigned long long* a = p.lbest + (size_t)q * p.lbest_stride;
    for (int b = lane; b < G; b += 32) stage[b] = a[b];
    __syncwarp();
    unsigned long long found = 0ull;
    for (int b = lane; b < G; b += 32) {
        const unsigned long long v = stage[b];
        if (v == 0ull) continue;
        int above = 0;
        for (int j = 0; j < G; j++) above += (stage[j] > v) ? 1 : 0;
        if (above == p.r_local - 1) found = v;      // words are unique: at most one lane hits
    }
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, found, d);
        found = o > found ? o : found;
    }
    __syncwarp();
    return found;
}

// ---- epilogue shared by the TMA and the generic kernel -----------------------
// Called by the kSelThreads consumer threads after the CTA's last tile.  big_buf/big_cap: a
// larger buffer the last CTA may merge in (the TMA kernel lends its idle stage ring), or nullptr.
template <int NQ>
__device__ __forceinline__ void scan_epilogue(const ScanParams& p, SelState* sel, int* flag, int* pref, int tid,
                                              uint64_t* big_buf, int big_cap) {
    const int k = p.topk;
    // 1. publish this CTA's survivors: the buffered words that reach the bound.  With the grid-wide
    // bound in place that is a handful of words, so the usual case needs no sort at all: every thread
    // copies its qualifying words out unsorted (the last CTA orders everything anyway) and the count
    // decides afterwards whether they fitted.
    for (int q = 0; q < NQ; q++) {
        uint64_t* dst = p.part + ((size_t)blockIdx.x * kMaxNQ + q) * k;
        bool fast = sel[q].m_local == 1;      // larger m publishes its m-th best from the sorted buffer
        if (fast) {
            const int n = *sel[q].cnt;
            const unsigned long long bound = *sel[q].tau;
            if (tid == 0) *flag = 0;
            named_bar_sync(kSelBarrier, kSelThreads);
            for (int i = tid; i < n; i += kSelThreads) {
                const uint64_t c = sel[q].buf[i];
                if (c >= bound) {
                    const int pos = atomicAdd(flag, 1);
                    if (pos < k) dst[pos] = c;
                }
            }
            named_bar_sync(kSelBarrier, kSelThreads);
            const int total = *((volatile int*)flag);
            named_bar_sync(kSelBarrier, kSelThreads);   // everyone has read the count before the next query resets it
            if (total <= k) {
                if (tid == 0) p.part_cnt[blockIdx.x * kMaxNQ + q] = total;
            } else {
                fast = false;
            }
        }
        if (!fast) {
            sel_prune(sel[q], tid, true);
            int m = *sel[q].cnt;
            unsigned long long bound = *sel[q].tau;   // includes the grid-wide bounds seen so far
            // survivors are sorted descending, so those >= bound form a prefix
            for (int i = tid; i < m; i += kSelThreads) {
                uint64_t c = sel[q].buf[i];
                if (c >= bound) dst[i] = c;
            }
            if (tid == 0) {
                int lo = 0, hi = m;   // first index with buf[i] < bound
                while (lo < hi) { int mid = (lo + hi) >> 1; if (sel[q].buf[mid] >= bound) lo = mid + 1; else hi = mid; }
                p.part_cnt[blockIdx.x * kMaxNQ + q] = lo;
            }
        }
    }
    __threadfence();
    named_bar_sync(kSelBarrier, kSelThreads);
    if (tid == 0) {
        unsigned int t = atomicAdd(p.ticket, 1u);
        *flag = (t == gridDim.x - 1) ? 1 : 0;
    }
    named_bar_sync(kSelBarrier, kSelThreads);
    if (*flag == 0) return;
    __threadfence();
    unsigned long long* lstamp = p.dbg_times ? p.dbg_times + (size_t)kMaxGrid * 4 : nullptr;   // development: the last CTA's phases
    if (lstamp && tid == 0) lstamp[0] = globaltimer_ns();

    // 2. last CTA: merge every CTA's published survivors, its own included (they are in `part` like
    // everybody's).  Its tau is a valid lower bound of the k-th best and may BE a row of the answer
    // (its own k-th best, or another CTA's row that travelled through gtau / lbest): admit words
    // >= tau, i.e. > tau - 1.  Every row is published by exactly one CTA, so nothing is offered twice.
    if (tid < NQ) {
        unsigned long long t = *sel[tid].tau;
        *sel[tid].tau = t ? t - 1 : 0ull;
        *sel[tid].cnt = 0;
    }
    named_bar_sync(kSelBarrier, kSelThreads);
    const int grid = (int)gridDim.x;
    for (int q = 0; q < NQ; q++) {
        // Flatten the other CTAs' lists: pref[b] = number of survivors in CTAs < b.  All loads of
        // one round are independent, so the merge costs a few L2 round trips, not one per CTA.
        for (int b = tid; b < grid; b += kSelThreads)
            pref[b + 1] = __ldcg(p.part_cnt + b * kMaxNQ + q);
        if (tid == 0) pref[0] = 0;
        named_bar_sync(kSelBarrier, kSelThreads);
        if (tid < 32) {   // inclusive scan of pref[1..grid] by warp 0
            const int per = (grid + 31) / 32;
            int lo = 1 + tid * per, hi = lo + per;
            if (hi > grid + 1) hi = grid + 1;
            int sum = 0;
            for (int i = lo; i < hi; i++) sum += pref[i];
            int incl = sum;
            for (int d = 1; d < 32; d <<= 1) {
                int v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                if (tid >= d) incl += v;
            }
            int run = incl - sum;
            for (int i = lo; i < hi; i++) { run += pref[i]; pref[i] = run; }
        }
        named_bar_sync(kSelBarrier, kSelThreads);
        const int total = pref[grid];
        if (lstamp && tid == 0 && q == 0) { lstamp[1] = globaltimer_ns(); lstamp[5] = (unsigned long long)total; }
        // merge buffer: everything in ONE round in the lent buffer when it fits, else rounds in the
        // query's own buffer (a huge survivor count means the bounds did not bite; rounds keep the
        // sorts small)
        SelState ms = sel[q];
        ms.gtau = nullptr; ms.lbest = nullptr; ms.lmax = nullptr;
        if (big_buf && big_cap >= 2 * sel[q].cap && total <= big_cap / 4) {
            ms.buf = big_buf; ms.cap = big_cap;
            ms.room = total > 0 ? total : 1;      // everything fits: one round, no prune before the final one
        }
        // Many survivors mean the CTAs finished before a grid-wide bound had formed (a store of one or two tiles
        // per CTA): fold in the bound their final bests give now, so that the offers below are a handful.
        if (p.lbest && total > kSelThreads && ms.buf == big_buf && big_cap / 2 >= grid) {
            if (tid < 32) {
                const unsigned long long v = lbest_rth(p, q, tid, (unsigned long long*)(big_buf + big_cap / 2));
                if (tid == 0 && v > 0ull && v - 1 > *ms.tau) *ms.tau = v - 1;   // v may BE a row of the answer: admit >= v
            }
            named_bar_sync(kSelBarrier, kSelThreads);
        }
        const int room = ms.room;
        for (int base = 0; base < total; base += room) {
            const int stop = base + room < total ? base + room : total;
            // four words per thread and trip: their loads are issued together, so a list of a few hundred words (a
            // store of one tile per CTA, where no grid-wide bound has formed yet) costs one L2 round trip, not three
            for (int i0 = base + tid; i0 < stop; i0 += 4 * kSelThreads) {
                uint64_t w[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int i = i0 + u * kSelThreads;
                    w[u] = 0ull;
                    if (i < stop) {
                        int lo = 0, hi = grid;   // largest b with pref[b] <= i
                        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (pref[mid] <= i) lo = mid; else hi = mid; }
                        w[u] = __ldcg(p.part + ((size_t)lo * kMaxNQ + q) * k + (i - pref[lo]));
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (i0 + u * kSelThreads < stop) sel_offer(ms, w[u]);
            }
            named_bar_sync(kSelBarrier, kSelThreads);
            sel_maybe_prune(ms, tid, false);
        }
        if (lstamp && tid == 0 && q == 0) lstamp[2] = globaltimer_ns();
        sel_prune(ms, tid, false);
        if (lstamp && tid == 0 && q == 0) lstamp[3] = globaltimer_ns();
        int m = *ms.cnt;
        const int o = scan_slot_of(p, q);
        for (int i = tid; i < k; i += kSelThreads) {
            size_t at = (size_t)o * k + i;
            if (i < m) {
                uint64_t c = ms.buf[i];
                uint32_t slot = cand_slot(c);
                const int64_t id = p.ids[slot];
                p.out_cand[at] = c;
                p.out_sims[at] = cand_score(c);
                p.out_slots[at] = slot;
                p.out_ids[at] = id;
                record_write(p.out_rec + (size_t)o * p.rec_bytes, p.rec_topk, i, id, cand_score(c));
            } else {
                p.out_cand[at] = 0ull;
                p.out_sims[at] = 0.0f;
                p.out_slots[at] = 0xFFFFFFFFu;
                p.out_ids[at] = -1;
            }
        }
        if (p.out_keys) {
            // the reply wants the winning rows (store.go:399-401): copy them here and save the host a kernel launch
            float* dst = p.out_keys + (size_t)o * k * p.dim;
            if ((p.dim & 3) == 0) {
                const int n4 = p.dim >> 2;
                for (int e = tid; e < m * n4; e += kSelThreads) {
                    const int i = e / n4, c4 = e - i * n4;
                    const float4* src = reinterpret_cast<const float4*>(p.keys + (size_t)cand_slot(ms.buf[i]) * p.dim);
                    reinterpret_cast<float4*>(dst + (size_t)i * p.dim)[c4] = __ldcs(src + c4);
                }
            } else {
                for (int e = tid; e < m * p.dim; e += kSelThreads) {
                    const int i = e / p.dim, cc = e - i * p.dim;
                    dst[(size_t)i * p.dim + cc] = p.keys[(size_t)cand_slot(ms.buf[i]) * p.dim + cc];
                }
            }
        }
        if (tid == 0) {
            p.out_n[o] = m;
            *(int*)(p.out_rec + (size_t)o * p.rec_bytes) = m;
            p.gtau[q] = 0ull;   // leave the scratch clean for the next launch on this context
        }
        if (p.lbest)
            for (int b = tid; b < grid; b += kSelThreads) p.lbest[(size_t)q * p.lbest_stride + b] = 0ull;
        named_bar_sync(kSelBarrier, kSelThreads);
    }
    if (tid == 0) { *p.ticket = 0u; *p.tile_counter = 0u; }
    if (lstamp && tid == 0) lstamp[4] = globaltimer_ns();
    if (p.xchg.world > 0) {
        // this launch's records are final (the loop above ended on a barrier): push them to the peers
        const size_t off = (size_t)p.out_first * p.rec_bytes;
        xchg_copy(p.xchg, p.out_rec + off, off, (size_t)NQ * p.rec_bytes, tid, kSelThreads);
        named_bar_sync(kSelBarrier, kSelThreads);
        xchg_raise(p.xchg, tid);
    }
}

template <int NQ>
__device__ __forceinline__ void sel_setup(const ScanParams& p, const ScanSmem& sm, SelState* sel, int tid) {
    for (int q = 0; q < NQ; q++) {
        sel[q].buf = sm.sel + (size_t)q * p.cap;
        sel[q].cnt = sm.cnt + q;
        sel[q].tau = sm.tau + q;
        sel[q].cap = p.cap;
        sel[q].k = p.topk;
        sel[q].room = p.topk > 2 * kRowsPerTile ? p.topk : 2 * kRowsPerTile;
        sel[q].gtau = p.gtau + q;
        sel[q].lbest = p.lbest ? p.lbest + (size_t)q * p.lbest_stride + blockIdx.x : nullptr;
        sel[q].m_local = p.m_local;
        sel[q].lmax = (sm.lmax && p.lbest && p.m_local == 1) ? sm.lmax + q : nullptr;
    }
    if (tid < NQ) { sm.cnt[tid] = 0; sm.tau[tid] = 0ull; }
}

// Offer one finished tile's scores and keep the buffers within capacity.
// ext_bound (threads < NQ): a grid-wide lower bound of the k-th best for "their" query.
// last_pub (threads < NQ): the lbest word this thread last published for its query.
template <int NQ, bool SPILL>
__device__ __forceinline__ void tile_offer(const ScanParams& p, SelState* sel, const float* score, uint32_t row,
                                           bool valid, int tid, unsigned long long ext_bound,
                                           unsigned long long& last_pub) {
    if constexpr (SPILL) {
        // big-k path: every row's candidate word goes to HBM, selection happens in a later pass
        if (valid) {
#pragma unroll
            for (int q = 0; q < NQ; q++) p.all_cand[(size_t)q * p.n_rows + row] = pack_cand(score[q], row);
        }
    } else {
        if (valid) {
#pragma unroll
            for (int q = 0; q < NQ; q++) sel_offer(sel[q], pack_cand(score[q], row));
        }
        named_bar_sync(kSelBarrier, kSelThreads);
        if (tid < NQ) {
            if (ext_bound > *sel[tid].tau) *sel[tid].tau = ext_bound;
            // keep this CTA's published "best so far" fresh (m == 1; larger m publishes at prunes)
            if (sel[tid].lmax && *sel[tid].lmax > last_pub) {
                last_pub = *sel[tid].lmax;
                *((volatile unsigned long long*)sel[tid].lbest) = last_pub;
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; q++) sel_maybe_prune(sel[q], tid, true);
    }
}

// ---- the TMA kernel -----------------------------------------------------------
template <int NQ, bool F64, bool SPILL>
__global__ void __launch_bounds__(kScanThreads, 1)
scan_tma_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ ScanParams p) {
    extern __shared__ uint8_t smem_raw[];
    const int dpad = p.n_chunks * kBoxCols;
    ScanSmem sm = scan_carve(smem_raw, p.n_stages, NQ, dpad, p.cap);
    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const uint32_t n_tiles = (p.n_rows + kRowsPerTile - 1) / kRowsPerTile;

    if (tid == 0) {
        for (int s = 0; s < p.n_stages; s++) {
            mbar_init(sm.full + s, 1);                     // producer's arrive.expect_tx
            mbar_init(sm.empty + s, kRowsPerTile / 32);    // one arrive per consumer warp
        }
        mbar_init(sm.qbar, 1);
        fence_mbar_init();
        *sm.done = 0;
    }
    if (tid < NQ) { sm.ext[tid] = 0ull; sm.lmax[tid] = 0ull; }
    // zero the padded query tail (columns >= dim) before the bulk copies land
    for (int i = tid; i < NQ * dpad; i += kScanThreads) {
        int c = i % dpad;
        if (c >= p.dim) sm.q[i] = 0.0f;
    }
    fence_proxy_async_smem();
    __syncthreads();

    if (warp == kRowsPerTile / 32) {
        // ===================== producer warp (one elected lane) =====================
        if ((tid & 31) == 0) {
            tma_prefetch_desc(&tmap);
            // TMA-stage the queries: one 1-D bulk copy per query
            mbar_arrive_expect_tx(sm.qbar, (uint32_t)(NQ * p.dim * 4));
            for (int q = 0; q < NQ; q++)
                bulk_load_1d(sm.q + (size_t)q * dpad, p.q + (size_t)scan_query_of(p, q) * p.dim, (uint32_t)(p.dim * 4),
                             sm.qbar);
            const uint64_t policy = p.evict_first ? l2_policy_evict_first() : l2_policy_evict_normal();
            int stage = 0;
            uint32_t phase = 0;
            // Dynamic tile scheduler: SMs do not pull from HBM at equal rates (measured: busy time
            // 712k..937k cycles under a static deal), so tiles are claimed from a grid-wide counter.
            // The first tile is static; the next claim is issued a whole tile ahead, so the atomic's
            // round trip is hidden behind ~9 us of TMA issue.
            uint32_t tile = blockIdx.x;
            while (tile < n_tiles) {
                const uint32_t next = gridDim.x + atomicAdd(p.tile_counter, 1u);
                for (int c = 0; c < p.n_chunks; c++) {
                    mbar_wait(sm.empty + stage, phase ^ 1u);
                    if (c == 0) sm.tile_id[stage] = tile;   // published by the arrive below
                    mbar_arrive_expect_tx(sm.full + stage, kStageBytes);
                    tma_load_2d(sm.stages + (size_t)stage * kStageBytes, &tmap, sm.full + stage, c * kBoxCols,
                                (int)(tile * kRowsPerTile), policy);
                    if (++stage == p.n_stages) { stage = 0; phase ^= 1u; }
                }
                tile = next;
            }
            // end marker: a stage with no bytes whose tile id says "done"
            mbar_wait(sm.empty + stage, phase ^ 1u);
            sm.tile_id[stage] = 0xFFFFFFFFu;
            mbar_arrive(sm.full + stage);
        }
        return;
    }
    if (warp == kRowsPerTile / 32 + 1) {
        // ===================== bound-refresher warp =====================
        // Polls the grid-wide bounds in HBM (gtau, and the min over CTAs of lbest) and keeps the
        // freshest value in shared memory, so the consumer threads never wait on a global load.
        if constexpr (!SPILL) {
            const int lane = tid & 31;
            if (p.refresh_ns <= 0) return;
            for (;;) {
                // one lane reads the flag and the warp follows it, so every lane leaves together
                int d = (lane == 0) ? *((volatile int*)sm.done) : 0;
                if (__shfl_sync(0xFFFFFFFFu, d, 0)) break;
                for (int q = 0; q < NQ; q++) {
                    unsigned long long v = p.lbest ? lbest_rth(p, q, lane, sm.lstage) : 0ull;
                    if (lane == 0) {
                        unsigned long long g = *((volatile unsigned long long*)(p.gtau + q));
                        *((volatile unsigned long long*)(sm.ext + q)) = g > v ? g : v;
                    }
                }
                __nanosleep((unsigned)p.refresh_ns);
            }
        }
        return;
    }

    // ========================= consumer threads: one row each =========================
    SelState sel[NQ];
    sel_setup<NQ>(p, sm, sel, tid);
    named_bar_sync(kSelBarrier, kSelThreads);

    double mag1[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) mag1[q] = F64 ? p.qinfo[scan_query_of(p, q)].mag1 : 0.0;

    if (p.dbg_times && tid == 0) p.dbg_times[blockIdx.x * 4 + 0] = globaltimer_ns();
    unsigned long long dbg_tiles = 0;
    mbar_wait(sm.qbar, 0);
    const uint32_t swz = (uint32_t)(tid & 7);
    const uint32_t row_off = (uint32_t)tid * 128u;
    int stage = 0;
    uint32_t phase = 0;
    unsigned long long last_pub = 0ull;
    for (;;) {
        Acc<F64> acc[NQ];
#pragma unroll
        for (int q = 0; q < NQ; q++) acc[q].zero();
        mbar_wait(sm.full + stage, phase);   // first chunk of the next tile, or the end marker
        const uint32_t tile = sm.tile_id[stage];
        if (tile == 0xFFFFFFFFu) break;
        const uint32_t row = tile * kRowsPerTile + (uint32_t)tid;
        const bool valid = row < p.n_rows;
        // issued now, consumed at the end of the tile: its latency hides behind the row walk
        const double kn = (F64 && valid) ? __ldg(p.knorm + row) : 1.0;
        for (int c = 0; c < p.n_chunks; c++) {
            if (c) mbar_wait(sm.full + stage, phase);
            const uint8_t* srow = sm.stages + (size_t)stage * kStageBytes + row_off;
            const float* qc = sm.q + c * kBoxCols;
            const int nvalid = p.dim - c * kBoxCols;  // >= 32 except in a ragged last chunk
            if (nvalid >= kBoxCols) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float4 kv = *reinterpret_cast<const float4*>(srow + (((uint32_t)j ^ swz) << 4));
#pragma unroll
                    for (int q = 0; q < NQ; q++) {
                        const float4 qv = *reinterpret_cast<const float4*>(qc + (size_t)q * dpad + j * 4);
                        acc[q].add(qv.x, kv.x);
                        acc[q].add(qv.y, kv.y);
                        acc[q].add(qv.z, kv.z);
                        acc[q].add(qv.w, kv.w);
                    }
                }
            } else {
                for (int j = 0; j * 4 < nvalid; j++) {  // dim % 4 == 0 on this path
                    const float4 kv = *reinterpret_cast<const float4*>(srow + (((uint32_t)j ^ swz) << 4));
#pragma unroll
                    for (int q = 0; q < NQ; q++) {
                        const float4 qv = *reinterpret_cast<const float4*>(qc + (size_t)q * dpad + j * 4);
                        acc[q].add(qv.x, kv.x);
                        acc[q].add(qv.y, kv.y);
                        acc[q].add(qv.z, kv.z);
                        acc[q].add(qv.w, kv.w);
                    }
                }
            }
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(sm.empty + stage);
            if (++stage == p.n_stages) { stage = 0; phase ^= 1u; }
        }
        float score[NQ];
#pragma unroll
        for (int q = 0; q < NQ; q++) score[q] = finish_score<F64>(acc[q], mag1[q], kn);
        unsigned long long ext_bound = 0ull;
        if (!SPILL && tid < NQ) ext_bound = *((volatile unsigned long long*)(sm.ext + tid));
        tile_offer<NQ, SPILL>(p, sel, score, row, valid, tid, ext_bound, last_pub);
        dbg_tiles++;
    }
    if (p.dbg_times && tid == 0) { p.dbg_times[blockIdx.x * 4 + 1] = globaltimer_ns(); p.dbg_times[blockIdx.x * 4 + 3] = dbg_tiles; }
    if (tid == 0) *((volatile int*)sm.done) = 1;   // the refresher warp may leave
    if constexpr (!SPILL) {
        // the stage ring is idle now (every TMA box has landed and been consumed): lend it to the merge
        int big_cap = 1;
        while ((size_t)big_cap * 2 * 8 <= (size_t)p.n_stages * kStageBytes) big_cap <<= 1;
        scan_epilogue<NQ>(p, sel, sm.flag, sm.pref, tid, (uint64_t*)sm.stages, big_cap);
        if (p.dbg_times && tid == 0) p.dbg_times[blockIdx.x * 4 + 2] = globaltimer_ns();
    }
}

// ---- the generic kernel: any dim (dim % 4 != 0, dim < 32, or a query too long
// for shared memory).  Same arithmetic and selection, one thread per row.  A thread that reads its own row
// straight from global memory makes every load instruction of its warp touch 32 lines (measured: 1.1 - 1.5 TB/s,
// and two queries per pass slower than two passes; one CTA of four warps per SM has nothing in flight to hide the
// latency either), so a warp fetches its 32 rows 32 columns at a time -- lane l copies column c0 + l of every row:
// one or two lines per instruction -- into a padded shared-memory tile (144-byte rows: each lane's 128-bit loads of
// its own row are conflict-free), with the same columns of the queries beside it.  The next step's loads (across
// tile boundaries too) are issued before this step's adds, and kGenCtasPerSm CTAs share an SM. -------
constexpr int kGenRowFloats = 36;              // staged row: 32 columns + 4 of padding
constexpr int kGenCtasPerSm = 3;               // resident CTAs (of four warps) per SM: what keeps enough loads in flight
__host__ __device__ inline size_t scan_generic_stage_floats(int nq) { return (size_t)32 * kGenRowFloats + (size_t)nq * 32; }
template <int NQ, bool F64, bool SPILL>
__global__ void __launch_bounds__(kSelThreads, 4)
scan_generic_kernel(const __grid_constant__ ScanParams p) {
    extern __shared__ uint8_t smem_raw[];
    ScanSmem sm;  // no stages, no staged query: only the selection state
    uint8_t* base = smem_raw + ((16u - (smem_u32(smem_raw) & 15u)) & 15u);
    sm.sel = (uint64_t*)base;
    sm.tau = (unsigned long long*)(base + (size_t)NQ * p.cap * 8);
    sm.cnt = (int*)(sm.tau + NQ);
    sm.flag = sm.cnt + NQ;
    sm.pref = sm.flag + 1;
    sm.lmax = nullptr;   // the generic kernel publishes lbest at prunes only
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    constexpr uint32_t kStageFloats = (uint32_t)(32 * kGenRowFloats + NQ * 32);     // a warp's tile, then its query columns
    float* ring = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(sm.pref + kMaxGrid + 1) + 15) & ~(uintptr_t)15) +
                  (size_t)wrp * kStageFloats;
    SelState sel[NQ];
    sel_setup<NQ>(p, sm, sel, tid);
    named_bar_sync(kSelBarrier, kSelThreads);
    double mag1[NQ];
    size_t qoff[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) {
        mag1[q] = F64 ? p.qinfo[scan_query_of(p, q)].mag1 : 0.0;
        qoff[q] = (size_t)scan_query_of(p, q) * p.dim;
    }
    const uint32_t n_tiles = (p.n_rows + kRowsPerTile - 1) / kRowsPerTile;
    const uint32_t n_chunks = ((uint32_t)p.dim + 31u) / 32u;
    const uint32_t my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    const uint32_t total = my_tiles * n_chunks;                                     // 32-column steps this warp walks, over all its tiles
    // producer side: the NEXT step's loads are issued before this step's adds and wait in registers meanwhile (4-byte
    // cp.async copies were measured at a third of this: ~19 cycles of the load/store unit each)
    uint32_t i_tile = blockIdx.x, i_chunk = 0;
    float nxt[32], nxq[NQ];
    auto fetch = [&](uint32_t k) {
        if (k < total) {
            const int c0 = (int)i_chunk * 32;
            const uint32_t row0 = i_tile * kRowsPerTile + (uint32_t)(wrp * 32);
            const int nr = row0 < p.n_rows ? (int)(p.n_rows - row0 < 32u ? p.n_rows - row0 : 32u) : 0;
            const bool on = c0 + lane < p.dim;
            const float* src = p.keys + (size_t)row0 * p.dim + c0 + lane;
#pragma unroll
            for (int r = 0; r < 32; r++) {
                const float* a = src + (size_t)r * p.dim;
                nxt[r] = (on && r < nr) ? (p.evict_first ? __ldcs(a) : __ldcg(a)) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < NQ; q++) nxq[q] = on ? __ldg(p.q + qoff[q] + c0 + lane) : 0.0f;
            if (++i_chunk == n_chunks) { i_chunk = 0; i_tile += gridDim.x; }
        }
    };
    fetch(0);
    unsigned long long last_pub = 0ull;
    uint32_t tile = blockIdx.x, chunk = 0;
    Acc<F64> acc[NQ];
    unsigned long long g_seen = 0ull;
    for (uint32_t k = 0; k < total; k++) {
        {
            float* st = ring;
#pragma unroll
            for (int r = 0; r < 32; r++) st[r * kGenRowFloats + lane] = nxt[r];
#pragma unroll
            for (int q = 0; q < NQ; q++) st[32 * kGenRowFloats + q * 32 + lane] = nxq[q];
        }
        __syncwarp();
        fetch(k + 1);                                                               // in flight while this step's adds run
        if (chunk == 0) {
#pragma unroll
            for (int q = 0; q < NQ; q++) acc[q].zero();
            g_seen = 0ull;
            if (!SPILL && tid < NQ) g_seen = *((volatile unsigned long long*)(p.gtau + tid));
        }
        const int w = p.dim - (int)chunk * 32 < 32 ? p.dim - (int)chunk * 32 : 32;
        const float* st = ring;
        const float4* kr4 = reinterpret_cast<const float4*>(st + lane * kGenRowFloats);
        const float* sq = st + 32 * kGenRowFloats;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (4 * j < w) {                                                        // (warp-uniform); the adds stay in index order
                const float4 kv = kr4[j];
#pragma unroll
                for (int q = 0; q < NQ; q++) {
                    const float4 qq = reinterpret_cast<const float4*>(sq + q * 32)[j];
                    acc[q].add(qq.x, kv.x);
                    if (4 * j + 1 < w) acc[q].add(qq.y, kv.y);
                    if (4 * j + 2 < w) acc[q].add(qq.z, kv.z);
                    if (4 * j + 3 < w) acc[q].add(qq.w, kv.w);
                }
            }
        }
        __syncwarp();                                                               // the tile may be overwritten
        if (++chunk == n_chunks) {
            chunk = 0;
            const uint32_t row = tile * kRowsPerTile + (uint32_t)tid;
            const bool valid = row < p.n_rows;
            float score[NQ];
            const double kn = (F64 && valid) ? p.knorm[row] : 1.0;
#pragma unroll
            for (int q = 0; q < NQ; q++) score[q] = finish_score<F64>(acc[q], mag1[q], kn);
            tile_offer<NQ, SPILL>(p, sel, score, row, valid, tid, g_seen, last_pub);
            tile += gridDim.x;
        }
    }
    if constexpr (!SPILL) scan_epilogue<NQ>(p, sel, sm.flag, sm.pref, tid, nullptr, 0);
}
__host__ inline size_t scan_generic_smem_bytes(int nq, int cap) {
    return (size_t)nq * cap * 8 + (size_t)nq * 8 + (size_t)nq * 4 + 4 + (size_t)(kMaxGrid + 1) * 4 + 32 +
           16 + (size_t)(kSelThreads / 32) * scan_generic_stage_floats(nq) * 4;
}

// ---- big-k path, pass 2: decode the first k of the (already sorted) spill -------
__global__ void decode_sorted_kernel(const uint64_t* sorted, uint32_t n_rows, int k, const int64_t* ids, int o,
                                     uint64_t* out_cand, int64_t* out_ids, float* out_sims, uint32_t* out_slots,
                                     int* out_n, uint8_t* out_rec, int rec_bytes, int rec_topk) {
    const int m = (uint32_t)k < n_rows ? k : (int)n_rows;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k; i += gridDim.x * blockDim.x) {
        size_t at = (size_t)o * k + i;
        if (i < m) {
            uint64_t c = sorted[i];
            uint32_t slot = cand_slot(c);
            const int64_t id = ids[slot];
            out_cand[at] = c;
            out_sims[at] = cand_score(c);
            out_slots[at] = slot;
            out_ids[at] = id;
            record_write(out_rec + (size_t)o * rec_bytes, rec_topk, i, id, cand_score(c));
        } else {
            out_cand[at] = 0ull;
            out_sims[at] = 0.0f;
            out_slots[at] = 0xFFFFFFFFu;
            out_ids[at] = -1;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { out_n[o] = m; *(int*)(out_rec + (size_t)o * rec_bytes) = m; }
}

}  // namespace tfp
