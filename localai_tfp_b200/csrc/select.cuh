// select.cuh -- fused top-k selection that lives inside the scan kernel.
//
// Replaces the reference's container/heap priority queue
// (backend/go/stores/store.go:341-371, push/pop per key at :380-388) with a
// threshold-filtered candidate buffer in shared memory:
//   * every row's score becomes one 64-bit candidate word (common.cuh);
//   * a row is appended only if it beats the current k-th best (`tau`);
//   * when the buffer is nearly full the CTA sorts it (bitonic, descending),
//     keeps the best k and raises tau;
//   * tau is also raised from two grid-wide bounds kept in HBM, both lower bounds
//     of the global k-th best, so rows below them can never be in the answer:
//       gtau  = max over CTAs of "my k-th best" (atomicMax);
//       lbest = min over CTAs of "my m-th best", m = ceil(k / #CTAs): every CTA
//               holds m rows at least that good, so #CTAs * m >= k rows are.
//     The second one is what keeps the per-CTA survivor lists short when k is
//     large.  Both only drop non-winners, so the result is deterministic.
// The selection costs nothing in HBM traffic: the key stream is read once.
#pragma once
#include "common.cuh"

namespace tfp {

constexpr int kRowsPerTile = 128;           // rows per tile == consumer threads per CTA
constexpr int kSelThreads = kRowsPerTile;   // threads that run the selection
constexpr int kSelBarrier = 1;              // named barrier id used by the consumer warps

struct SelState {
    uint64_t* buf;           // [cap] candidate words (shared)
    int* cnt;                // entries in buf (shared)
    unsigned long long* tau; // current admission threshold (shared)
    int cap;                 // power of two
    int k;
    int room;                // pushes that may happen between two capacity checks
    unsigned long long* gtau;   // grid-wide max of local k-th bests (HBM), or nullptr
    unsigned long long* lbest;  // this CTA's slot in the per-CTA "m-th best" array (HBM), or nullptr
    int m_local;                // m = ceil(k / #CTAs)
    unsigned long long* lmax;   // shared: best word this CTA has admitted (tracked when m == 1), or nullptr
};

// Bitonic sort, descending, of buf[0..P) with P a power of two, by kSelThreads threads.
__device__ __forceinline__ void sel_sort_desc(uint64_t* buf, int P, int tid) {
    for (int k2 = 2; k2 <= P; k2 <<= 1) {
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (P >> 1); i += kSelThreads) {
                int lo = ((i / j) * (j << 1)) + (i % j);
                int hi = lo + j;
                uint64_t x = buf[lo], y = buf[hi];
                bool desc = ((lo & k2) == 0);
                if ((x < y) == desc && x != y) {
                    buf[lo] = y;
                    buf[hi] = x;
                }
            }
            named_bar_sync(kSelBarrier, kSelThreads);
        }
    }
}

// Rank sort for short lists: every word's place is the number of words larger than it (the words are
// unique: a row appears once), so the best k land in order with two barriers instead of one per bitonic
// stage -- the per-stage barrier is what a sort of a few dozen words by 128 threads spends its time on.
template <int SLOTS>
__device__ __forceinline__ void sel_rank_sort(const SelState& s, int n, int tid) {
    uint64_t v[SLOTS];
    int rk[SLOTS];
#pragma unroll
    for (int e = 0; e < SLOTS; e++) {
        const int i = tid + e * kSelThreads;
        v[e] = i < n ? s.buf[i] : ~0ull;      // a padding slot outranks everything: it never counts, never lands
        rk[e] = 0;
    }
    for (int j = 0; j < n; j++) {
        const uint64_t w = s.buf[j];          // same address for the whole warp: a broadcast
#pragma unroll
        for (int e = 0; e < SLOTS; e++) rk[e] += (w > v[e]) ? 1 : 0;
    }
    named_bar_sync(kSelBarrier, kSelThreads);  // every read is done before the first word moves
#pragma unroll
    for (int e = 0; e < SLOTS; e++)
        if (tid + e * kSelThreads < n && rk[e] < s.k) s.buf[rk[e]] = v[e];
    named_bar_sync(kSelBarrier, kSelThreads);
}

// Keep the best k of buf[0..cnt); raise tau; publish this CTA's bounds.
// Must be called by all kSelThreads threads with a uniform view of *cnt.
__device__ __forceinline__ void sel_prune(const SelState& s, int tid, bool publish) {
    int n = *s.cnt;                            // (thread 0 rewrites it only after the barriers of the sort)
    if (n <= kSelThreads) {
        sel_rank_sort<1>(s, n, tid);
    } else if (n <= 4 * kSelThreads) {
        sel_rank_sort<4>(s, n, tid);
    } else {
        int P = 2;
        while (P < n) P <<= 1;
        for (int i = n + tid; i < P; i += kSelThreads) s.buf[i] = 0ull;
        named_bar_sync(kSelBarrier, kSelThreads);
        sel_sort_desc(s.buf, P, tid);
    }
    if (tid == 0) {
        int m = n < s.k ? n : s.k;
        unsigned long long t = *s.tau;
        if (m == s.k) {
            unsigned long long kth = s.buf[s.k - 1];
            if (kth > t) t = kth;
            if (publish && s.gtau) {
                unsigned long long g = atomicMax(s.gtau, kth);
                if (g > t) t = g;
            }
        }
        if (publish && s.lbest && n >= s.m_local)
            *((volatile unsigned long long*)s.lbest) = s.buf[s.m_local - 1];
        *s.tau = t;
        *s.cnt = m;
    }
    named_bar_sync(kSelBarrier, kSelThreads);
}

// One thread offers one candidate.  Capacity is guaranteed by sel_maybe_prune.
__device__ __forceinline__ void sel_offer(const SelState& s, uint64_t cand) {
    if (cand > *((volatile unsigned long long*)s.tau)) {
        int pos = atomicAdd(s.cnt, 1);
        s.buf[pos] = cand;
        // m == 1: the best admitted word is this CTA's lbest.  Admissions are rare, so this costs
        // nothing per row; a row that was not admitted is below tau and cannot matter.
        if (s.lmax && cand > *((volatile unsigned long long*)s.lmax)) atomicMax(s.lmax, cand);
    }
}

// Call after a round of offers, with a barrier already between the offers and
// this call.  Uniform across the consumer threads.
__device__ __forceinline__ void sel_maybe_prune(const SelState& s, int tid, bool publish) {
    int n = *((volatile int*)s.cnt);
    named_bar_sync(kSelBarrier, kSelThreads);  // everyone has read cnt before anyone changes it
    if (n > s.cap - s.room) sel_prune(s, tid, publish);
}

}  // namespace tfp
