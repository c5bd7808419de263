// ingest.cuh -- device side of StoresSet / StoresDelete / StoresGet and the reply gather.
//
// The reference keeps keys sorted and merges/rebuilds two Go slices on every
// mutation (backend/go/stores/store.go:150-195, :219-246).  Here rows live
// unsorted in HBM; exact-match lookups go through a host hash index whose
// hashes are computed on the device, candidates are confirmed on the device
// with Go's key equality (cmp.Compare: -0 == +0, NaN == NaN), new rows are
// appended by a scatter kernel and deletions fill holes with tail rows.
#pragma once
#include "common.cuh"

namespace tfp {

// canonical bit pattern under Go's cmp.Compare equality
__host__ __device__ __forceinline__ uint32_t canon_bits(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4);
#endif
    if ((u & 0x7FFFFFFFu) > 0x7F800000u) return 0x7FC00000u;  // every NaN is the same key element
    if (u == 0x80000000u) return 0u;                          // -0 == +0
    return u;
}
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// Row hash = mix(sum_i mix(i, bits_i)): position-dependent, order-free to evaluate.
__host__ __device__ __forceinline__ uint64_t hash_term(uint32_t i, float f) {
    return mix64(((uint64_t)i << 32) | (uint64_t)canon_bits(f));
}
__host__ __device__ __forceinline__ uint64_t hash_finish(uint64_t sum, int dim) {
    return mix64(sum ^ ((uint64_t)(uint32_t)dim * 0xD6E8FEB86659FD93ull));
}

// Per staged row: hash, isNormalized (store.go:313-324), and sqrt(mag2) in the reference's arithmetic
// (store.go:420: fp64 sum of fp32-rounded squares, index order).  The sums of one row are sequential on purpose --
// they must round exactly as the Go loops do -- so one THREAD owns one row; but the rows are read the way HBM
// likes: a CTA of 64 threads owns 64 rows and walks them in [64 rows x 32 columns] tiles that its warps load
// with coalesced 128-byte row segments into shared memory (row stride 33 floats: thread t then reads its own row
// bank-conflict free).  The first version let every thread stride through its own row in global memory:
// 54 GB/s at dim 1536, under 1 % of HBM.  64 rows, not 128: the tile (8.4 KB) must fit into the shared memory a
// persistent scan CTA leaves free on its SM (~13 KB), or every StoresSet waits for the running finds to end.
constexpr int kStatRows = 64;
constexpr int kStatCols = 32;
__global__ void __launch_bounds__(kStatRows)
row_stats_kernel(const float* __restrict__ rows, int64_t n, int dim, uint64_t* __restrict__ hash,
                 double* __restrict__ knorm, uint8_t* __restrict__ is_norm) {
    __shared__ float tile[kStatRows][kStatCols + 1];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int64_t base = (int64_t)blockIdx.x * kStatRows;
    const int64_t r = base + t;
    double sum = 0.0, mag2 = 0.0;
    uint64_t h = 0;
    for (int c0 = 0; c0 < dim; c0 += kStatCols) {
        const int w = dim - c0 < kStatCols ? dim - c0 : kStatCols;
        for (int rr = warp; rr < kStatRows; rr += kStatRows / 32) {
            const int64_t row = base + rr;
            if (row < n && lane < w) tile[rr][lane] = __ldcs(rows + (size_t)row * dim + c0 + lane);
        }
        __syncthreads();
        if (r < n) {
#pragma unroll 8
            for (int j = 0; j < w; j++) {
                const float v = tile[t][j];
                const double v64 = (double)v;
                sum = __dadd_rn(sum, __dmul_rn(v64, v64));           // sum += v64*v64 (exact product)
                mag2 = __dadd_rn(mag2, (double)__fmul_rn(v, v));     // mag2 += float64(k2[i]*k2[i])
                h += hash_term((uint32_t)(c0 + j), v);
            }
        }
        __syncthreads();
    }
    if (r >= n) return;
    const double s = __dsqrt_rn(sum);
    hash[r] = hash_finish(h, dim);
    knorm[r] = __dsqrt_rn(mag2);
    is_norm[r] = (s >= 0.99 && s <= 1.01) ? 1 : 0;
}

// One warp per (staged row, stored slot) pair: are they the same key under
// Go's slices.Compare(...) == 0  (store.go:48-52)?
__global__ void verify_pairs_kernel(const float* __restrict__ keys, int dim, const float* __restrict__ rows,
                                    const uint32_t* __restrict__ pair_row, const uint32_t* __restrict__ pair_slot,
                                    int64_t n_pairs, uint8_t* __restrict__ eq) {
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (w >= n_pairs) return;
    const float* a = rows + (size_t)pair_row[w] * dim;
    const float* b = keys + (size_t)pair_slot[w] * dim;
    bool same = true;
    for (int i = lane; i < dim; i += 32) same = same && (canon_bits(a[i]) == canon_bits(b[i]));
    same = __all_sync(0xFFFFFFFFu, same);
    if (lane == 0) eq[w] = same ? 1 : 0;
}

// Append / overwrite: staged row src[i] -> slot dst[i], with its id and norm.
__global__ void scatter_rows_kernel(const float* __restrict__ rows, const double* __restrict__ rows_knorm,
                                    const uint32_t* __restrict__ src, const uint32_t* __restrict__ dst,
                                    const int64_t* __restrict__ new_ids, int64_t m, int dim,
                                    float* __restrict__ keys, int64_t* __restrict__ ids, double* __restrict__ knorm) {
    int64_t i = blockIdx.x;
    if (i >= m) return;
    const float* a = rows + (size_t)src[i] * dim;
    float* b = keys + (size_t)dst[i] * dim;
    for (int c = threadIdx.x; c < dim; c += blockDim.x) b[c] = a[c];
    if (threadIdx.x == 0) {
        ids[dst[i]] = new_ids[i];
        knorm[dst[i]] = rows_knorm[src[i]];
    }
}

// Delete compaction: move tail row src[i] into hole dst[i].  All src >= new_n > all dst.
__global__ void move_rows_kernel(float* __restrict__ keys, int64_t* __restrict__ ids, double* __restrict__ knorm,
                                 const uint32_t* __restrict__ src, const uint32_t* __restrict__ dst, int64_t m,
                                 int dim) {
    int64_t i = blockIdx.x;
    if (i >= m) return;
    const float* a = keys + (size_t)src[i] * dim;
    float* b = keys + (size_t)dst[i] * dim;
    for (int c = threadIdx.x; c < dim; c += blockDim.x) b[c] = a[c];
    if (threadIdx.x == 0) {
        ids[dst[i]] = ids[src[i]];
        knorm[dst[i]] = knorm[src[i]];
    }
}

// Reply gather: StoresFindResult.Keys wants the winning rows (store.go:399-401).
__global__ void gather_rows_kernel(const float* __restrict__ keys, int dim, const uint32_t* __restrict__ slots,
                                   const int* __restrict__ out_n, int topk, float* __restrict__ out) {
    int qi = blockIdx.y;
    int i = blockIdx.x;
    if (i >= out_n[qi]) return;
    const float* a = keys + (size_t)slots[(size_t)qi * topk + i] * dim;
    float* b = out + ((size_t)qi * topk + i) * dim;
    for (int c = threadIdx.x; c < dim; c += blockDim.x) b[c] = a[c];
}

// Cross-shard merge: one CTA per query, bitonic sort of
// n_shards x topk words (sim order, then lower shard, then position).  Input is the
// gathered block of packed records: shard sh's nq records (rec_bytes each) start at sh * shard_stride
// bytes.  Read with ld.cg: with the peer exchange the records were written by other GPUs over NVLink.
// wait_words != nullptr (peer exchange): the kernel itself waits until every peer's sequence word has
// reached wait_seq -- the peers' scan kernels raise them with st.release.sys after their records have
// landed -- so a sharded find is the scan plus THIS one launch, with no stream wait operations in
// between.  The CTA is small (no dynamic shared memory to speak of), so it co-resides with the scan
// CTAs of the other lanes while it waits.
__global__ void merge_records_kernel(const uint8_t* g, int rec_bytes, size_t shard_stride, int n_shards, int nq, int topk, int P,
                                     float* __restrict__ o_sims, int64_t* __restrict__ o_ids,
                                     int* __restrict__ o_shard, int* __restrict__ o_index, int* __restrict__ o_n,
                                     const unsigned int* wait_words, int wait_me, unsigned int wait_seq) {
    extern __shared__ uint64_t mbuf[];
    const int q = blockIdx.x;
    const int total = n_shards * topk;
    if (wait_words) {
        if ((int)threadIdx.x < n_shards && (int)threadIdx.x != wait_me) {
            const unsigned int* w = wait_words + threadIdx.x;
            unsigned int v;
            unsigned long long t0 = 0ull;
            for (;;) {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(w) : "memory");
                if ((int)(v - wait_seq) >= 0) break;
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                if (t0 == 0ull) t0 = now;
                else if (now - t0 > 10000000000ull) {   // 10 s: a peer never ran this find (the ranks' call sequences differ)
                    printf("tfp: sharded find timed out waiting for the records of rank %d (seq %u, saw %u)\n",
                           (int)threadIdx.x, wait_seq, v);
                    __trap();
                }
                __nanosleep(200);
            }
        }
        __syncthreads();
    }
    auto rec = [&](int sh) { return g + (size_t)sh * shard_stride + (size_t)q * rec_bytes; };
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
        uint64_t w = 0ull;
        if (i < total) {
            int sh = i / topk, j = i - sh * topk;
            const uint8_t* r = rec(sh);
            if (j < __ldcg((const int*)r)) {
                float s = __ldcg((const float*)(r + 8 + (size_t)topk * 8) + j);
                // low word: earlier (shard, position) ranks higher among equal scores
                w = ((uint64_t)score_to_order(s) << 32) | (uint64_t)(0xFFFFFFFFu - (uint32_t)i);
            }
        }
        mbuf[i] = w;
    }
    __syncthreads();
    for (int k2 = 2; k2 <= P; k2 <<= 1) {
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < (P >> 1); i += blockDim.x) {
                int lo = ((i / j) * (j << 1)) + (i % j), hi = lo + j;
                uint64_t x = mbuf[lo], y = mbuf[hi];
                bool desc = ((lo & k2) == 0);
                if ((x < y) == desc && x != y) { mbuf[lo] = y; mbuf[hi] = x; }
            }
            __syncthreads();
        }
    }
    int valid = 0;
    for (int sh = 0; sh < n_shards; sh++) {
        int c = __ldcg((const int*)rec(sh));
        valid += c < topk ? c : topk;
    }
    const int m = valid < topk ? valid : topk;
    for (int i = threadIdx.x; i < topk; i += blockDim.x) {
        size_t at = (size_t)q * topk + i;
        if (i < m) {
            uint32_t pos = 0xFFFFFFFFu - (uint32_t)mbuf[i];
            int sh = (int)(pos / (uint32_t)topk), j = (int)(pos - (uint32_t)sh * topk);
            const uint8_t* r = rec(sh);
            o_sims[at] = __ldcg((const float*)(r + 8 + (size_t)topk * 8) + j);
            o_ids[at] = __ldcg((const long long*)(r + 8) + j);
            o_shard[at] = sh;
            o_index[at] = j;
        } else {
            o_sims[at] = 0.0f;
            o_ids[at] = -1;
            o_shard[at] = -1;
            o_index[at] = -1;
        }
    }
    if (threadIdx.x == 0) o_n[q] = m;
}

}  // namespace tfp
