// mma_scan.cuh -- batched StoresFind on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// For batches of queries (3 or more: the measured break-even) the scan is a dense contraction S = K . Q^T
// (K: [rows x D] keys, Q: [nq x D] queries).  It replaces nq sequential runs of the
// reference loop (backend/go/stores/store.go:378-389 / :441-452) -- the reference has no
// batched path at all.
//
// Numerics: tensor cores only FILTER; finalize (below) RE-SCORES the surviving rows in the
// reference's exact arithmetic (index-order fp32 dot, or fp64-accumulated cosine), so the
// similarities returned are bit-identical to the Go code, exactly as on the single-query path.
// Two filters (template TERMS):
//   TERMS = 1  one TF32 MMA per K-step on the raw fp32 operands.  TF32 keeps 10 mantissa bits, so
//              each operand is off by < 2^-10 relative and |approx - exact| <= eps with
//              eps = (2^-9 + 2^-20 + 2 D 2^-23) |q| |k|: the operand term is a bound (Cauchy-Schwarz), the
//              D term covers the tensor core's own fp32 accumulation and the reference's fp32 rounding
//              with margin (worst case over what is known of that adder; see tfp_store.cu mma_pass_launch).
//              Every row within 2 eps of the k-th best approximate score is re-scored: a row outside that
//              band cannot be in the exact top-k while the bound holds.
//   TERMS = 3  each operand split into TF32 hi + lo (a = hi + lo to 2^-22), S ~= hi.hi + hi.lo +
//              lo.hi, error ~1e-6; the best k+48 candidates are re-scored and the query is re-run on
//              the exact kernel if the first row not re-scored is too close to call.  3x the MMAs
//              and twice the query traffic: shared-memory-bandwidth bound at half the tensor peak.
// Per query, candidates are the rows whose approximate score clears a lower bound (valid by counting) of the
// need-th best score (minus 2 eps).  A query whose candidate list overflowed, came out short, or whose band holds
// more rows than may be re-scored is flagged: the host sends the flagged queries of a batch through one 3xTF32
// pass together, and what that cannot settle through the exact scan kernel, 8 per pass -- all still on the GPU.
//
// Kernel anatomy (one persistent CTA per SM, 12 warps):
//   warp 0      TMA producer: per 16-column chunk one key box [128 rows x 64 B] (HBM, read once)
//               and the query box(es) [nq x 64 B] (L2-resident), 64-B swizzle, into a ring of
//               8 stages of 24 KiB (TERMS = 1) or 4 of 48 KiB (TERMS = 3) at nq = 256.  (Sharing one
//               query box between two key tiles was tried: it costs the second accumulator buffer,
//               and the lost MMA/epilogue overlap outweighed the saved L2 traffic.)
//   warps 8-11  converters: split the landed key tile into hi (in place) and lo (second tile),
//               fence.proxy.async so the tensor core sees the generic-proxy writes;
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma.kind::tf32 (M=128, N=nq, K=8),
//               TERMS per K-step, accumulating in one of two TMEM buffers; tcgen05.commit releases
//               the shared-memory stage and, after the last chunk, hands the accumulator over;
//   warps 4-7   epilogue: tcgen05.ld the finished accumulator (thread = key row, registers =
//               queries), scale by 1/|k| on the cosine path, compare against the per-query
//               threshold, append survivors to the query's candidate list in HBM;
//   warp 2      bound refresher: publishes this CTA's second-best score per query, computes the grid-wide
//               bound of "its" queries (the r-th largest second-best over the CTAs: r CTAs hold two rows
//               each at least that good) and keeps all thresholds fresh from the shared bounds.
// No score matrix is ever written to HBM.
#pragma once
#include "scan.cuh"

namespace tfp {

constexpr int kMmaThreads = 384;
constexpr int kMmaStages3 = 4;          // TERMS = 3: stages of 48 KiB at nq = 256
constexpr int kMmaStages1 = 8;          // TERMS = 1: stages of 24 KiB
constexpr int kMmaStagesMax = 24;        // small batches have small query boxes: more key tiles in flight
constexpr int kMmaHitQueue = 128;       // per epilogue warp (8 of them): hits staged in shared memory, flushed once per tile
constexpr int kMmaRescoreMax = 1024;    // TERMS = 1: rows re-scored per query at most (else exact re-run)
constexpr int kMmaBoxCols = 16;         // floats per row per stage (64 B, 64-byte swizzle): small stages, deep ring
constexpr int kMmaMaxNQ = 256;         // queries per pass == UMMA N (multiple of 16)
constexpr int kMmaCandCap = 16384;     // candidate words per query
constexpr int kMmaMargin = 48;         // candidates re-scored beyond k
constexpr int kMmaMaxTopK = 100;       // k + margin must stay within the 148 rows the bound guarantees
constexpr int kMmaRows = kRowsPerTile;                       // key rows per stage == one 128-row MMA tile
constexpr int kMmaATile = kMmaRows * kMmaBoxCols * 4;        // 8 KiB
constexpr int kMmaRowBytes = kMmaBoxCols * 4;                // 64

struct MmaParams {
    uint32_t n_rows;
    int dim, n_chunks;
    int nq, nqp;                 // queries in this pass, padded to a multiple of 16
    int f64;                     // cosine path: approximate score = dot / |k|
    int mode;                    // 0 = bound-only pass over the first tile of every CTA, 1 = full pass
    const double* knorm;
    unsigned int* lbest2;        // [grid x kMmaMaxNQ] orderable word of each CTA's SECOND best approximate score of this pass
    unsigned int* gbound;        // [kMmaMaxNQ] grid-wide lower bound of the need-th best score per query (only rises)
    int r_need;                  // r = ceil(need / 2): r CTAs with two rows each >= the bound make `need` rows
    uint64_t* cand;              // [nq x kMmaCandCap]
    unsigned int* cand_cnt;      // [nq]
    int refresh_ns;
    const float* eps;            // [nq] error bound of the approximate score (0 for TERMS = 3)
    int n_stages;                // ring depth (mma_stage_count)
    int box_cols;                // floats per row per stage: 16 (64-B swizzle) or 32 (128-B swizzle)
    int debug;                   // development only: 1 = skip the epilogue's per-element work, 2 = issue no MMAs
};

// ---- tcgen05 wrappers ---------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, TF32 inputs, fp32 accumulate
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// mbarrier arrives when every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// ---- CTA pair (cta_group::2): two SMs of one TPC run ONE M = 256 MMA; each holds its own 128 key rows
// (and their accumulator, in its own TMEM) and HALF of the query tile, which the tensor core reads from
// both shared memories -- so every SM pulls half the query bytes from L2 that a lone CTA would. ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory location in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// the same-offset mbarrier of BOTH CTAs of the pair arrives when the MMAs issued so far have retired
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// TMA load whose completion is signalled on an mbarrier that may live in the peer CTA (cluster address)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1,
                                                 uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
        "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
// One lane of a fully active warp (the rest of the warp keeps executing the same, warp-uniform code)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// K-major, 64-byte-swizzled operand tile: rows of 64 B, 8-row groups 512 B apart
// (cute::UMMA::SmemDescriptor: start>>4 | LBO=1<<16 | SBO=32<<32 | version=1<<46 | SWIZZLE_64B=4<<61)
__device__ __forceinline__ uint64_t umma_desc_sw64(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (32ull << 32) | (1ull << 46) | (4ull << 61);
}
// same with 128-byte rows / 128-byte swizzle: 8-row groups 1024 B apart, SWIZZLE_128B = 2
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, bool wide) {
    return wide ? umma_desc_sw128(smem_addr) : umma_desc_sw64(smem_addr);
}
// kind::tf32, fp32 accumulator, both operands K-major, M = 128, N = n
__host__ __device__ __forceinline__ uint32_t umma_idesc_tf32(int n, int m = 128) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
// 16-byte shared load that is re-issued every time (other warps update these words)
__device__ __forceinline__ float4 __uint4_as_float4(uint4 u) {
    return make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z), __uint_as_float(u.w));
}
__device__ __forceinline__ uint4 lds_volatile_v4(const unsigned int* p) {
    uint4 v;
    asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}
constexpr uint32_t kTf32Mask = 0xFFFFE000u;   // sign, exponent, 10 mantissa bits
__device__ __forceinline__ void tf32_split(float a, float& hi, float& lo) {
    hi = __uint_as_float(__float_as_uint(a) & kTf32Mask);
    lo = __uint_as_float(__float_as_uint(__fsub_rn(a, hi)) & kTf32Mask);   // a - hi is exact in fp32
}

// ---- query preparation: [nq x D] -> hi parts in rows [0, nqp), lo parts in rows [nqp, 2 nqp) -------
// Row r of the pass is staged query qmap[r] (a pass may serve any subset of the staged queries).
__global__ void qsplit_kernel(const float* __restrict__ q, const int* __restrict__ qmap, int nq, int nqp, int dim,
                              float* __restrict__ out) {
    const size_t total = (size_t)nqp * dim;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / dim);
        float hi = 0.0f, lo = 0.0f;
        if (r < nq) tf32_split(q[(size_t)qmap[r] * dim + (i - (size_t)r * dim)], hi, lo);
        out[i] = hi;
        out[total + i] = lo;
    }
}

// The single-TF32 filter moves 32 columns per stage (a 16 KiB key box: the per-stage cost of barriers,
// TMA issue and tcgen05.commit is then paid half as often, which is what small batches are bound by);
// the 3xTF32 filter, with four tiles per stage, stays at 16 columns to keep four stages.
__host__ __device__ inline int mma_box_cols(int terms) { return terms == 3 ? 16 : 32; }
// pair: a CTA of a cta_group::2 pair stages only its half of the query tile
__host__ __device__ inline size_t mma_stage_bytes(int nqp, int terms, bool pair = false) {
    const size_t row = (size_t)mma_box_cols(terms) * 4;
    if (pair) return (size_t)kMmaRows * row + (size_t)(nqp / 2) * row;
    return terms == 3 ? 2 * (size_t)kMmaRows * row + 2 * (size_t)nqp * row : (size_t)kMmaRows * row + (size_t)nqp * row;
}
__host__ __device__ inline size_t mma_fixed_smem_bytes() {
    return (3 * kMmaStagesMax + 4) * 8 + 16 + 5 * kMmaMaxNQ * 4 + kMaxGrid * 4 + 1024 + 8 * (size_t)kMmaHitQueue * 16 + 32;
}
// as many stages as fit (at least the pipeline's nominal depth at nq = 256, at most kMmaStagesMax)
__host__ __device__ inline int mma_stage_count(int nqp, int terms, size_t max_smem, bool pair = false) {
    const size_t room = max_smem > mma_fixed_smem_bytes() ? max_smem - mma_fixed_smem_bytes() : 0;
    int n = (int)(room / mma_stage_bytes(nqp, terms, pair));
    return n > kMmaStagesMax ? kMmaStagesMax : n;
}
__host__ __device__ inline size_t mma_smem_bytes(int nqp, int terms, int stages, bool pair = false) {
    return stages * mma_stage_bytes(nqp, terms, pair) + mma_fixed_smem_bytes();
}

// The r-th largest (with multiplicity) of G words spaced kMmaMaxNQ apart, by one warp; 0 = "nothing" and is
// skipped.  Returns 0 when fewer than r words are set.  `stage`: G words of shared scratch.
__device__ __forceinline__ unsigned int mma_rth_largest(const unsigned int* col, int G, int r, int lane, unsigned int* stage) {
    for (int b = lane; b < G; b += 32) stage[b] = ((const volatile unsigned int*)col)[(size_t)b * kMmaMaxNQ];
    __syncwarp();
    unsigned int found = 0u;
    for (int b = lane; b < G; b += 32) {
        const unsigned int v = stage[b];
        if (v == 0u || v <= found) continue;
        int ge = 0;
        for (int j = 0; j < G; j++) ge += (stage[j] >= v) ? 1 : 0;
        if (ge >= r) found = v;                           // at least r words reach v: v is a valid bound
    }
    for (int d = 16; d > 0; d >>= 1) {
        const unsigned int o = __shfl_xor_sync(0xFFFFFFFFu, found, d);
        found = o > found ? o : found;
    }
    __syncwarp();
    return found;
}
// Between the bound-only pass and the full pass: the first grid-wide bounds, one warp per query, so that
// the full pass starts with every threshold in place.
__global__ void mma_bound_kernel(const unsigned int* lbest2, int grid, int r, int nq, unsigned int* gbound) {
    __shared__ unsigned int stage[8][kMaxGrid];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + w;
    if (q >= nq) return;
    const unsigned int v = mma_rth_largest(lbest2 + q, grid, r, lane, stage[w]);
    if (lane == 0) gbound[q] = v;
}

// PAIR: launched as clusters of two CTAs (cta_group::2, TERMS = 1 only).  CTA r of a pair owns key tile
// blockIdx.x + i * gridDim.x like a lone CTA and stages query rows [r nqp/2, (r+1) nqp/2); the even CTA
// ("leader") issues every MMA for both.  Barriers: full[] lives on the leader and counts the bytes of both
// CTAs' TMA loads; empty[] / acc_full[] are signalled in both CTAs by one multicast commit; acc_empty[] on
// the leader collects the epilogue warps of both CTAs.  Both CTAs walk the same number of tiles (a tile
// index past the end is all padding), so the pair stays in step to the last barrier.
template <int TERMS, bool PAIR>
__global__ void __launch_bounds__(kMmaThreads, 1)
mma_scan_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_q, const MmaParams p) {
    static_assert(!(PAIR && TERMS == 3), "the CTA-pair variant exists for the single-TF32 filter only");
    const int kMmaStages = p.n_stages;
    const uint32_t crank = PAIR ? cluster_ctarank() : 0u;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int bc = p.box_cols;                                   // floats per row per stage
    const size_t a_bytes = (size_t)kMmaRows * bc * 4;
    const size_t b_bytes = (size_t)(PAIR ? p.nqp / 2 : p.nqp) * bc * 4;   // query bytes THIS CTA stages per chunk
    const bool wide = bc == 32;
    const size_t stage_bytes = mma_stage_bytes(p.nqp, TERMS, PAIR);
    // TERMS = 3: [a_hi | a_lo | b_hi | b_lo];  TERMS = 1: [a | b]
    auto a_hi = [&](int s) { return base + s * stage_bytes; };
    auto a_lo = [&](int s) { return base + s * stage_bytes + a_bytes; };
    auto b_hi = [&](int s) { return base + s * stage_bytes + (TERMS == 3 ? 2 : 1) * a_bytes; };
    auto b_lo = [&](int s) { return base + s * stage_bytes + 2 * a_bytes + b_bytes; };
    uint8_t* tail = base + kMmaStages * stage_bytes;
    uint64_t* full = (uint64_t*)tail;            // [stages] TMA landed
    uint64_t* conv = full + kMmaStagesMax;       // [stages] hi/lo tiles ready (TERMS = 3)
    uint64_t* empty = conv + kMmaStagesMax;      // [stages] stage consumed by the tensor core
    uint64_t* acc_full = empty + kMmaStagesMax;  // [2] accumulator complete
    uint64_t* acc_empty = acc_full + 2;          // [2] accumulator drained
    uint32_t* tmem_base_s = (uint32_t*)(acc_empty + 2);
    volatile int* flags = (volatile int*)(tmem_base_s + 1);   // [0] done, [1] thresholds ready
    unsigned int* tau = (unsigned int*)(tmem_base_s + 4);     // [kMmaMaxNQ] admission thresholds (orderable words)
    unsigned int* lb = tau + kMmaMaxNQ;                       // [kMmaMaxNQ] this CTA's best score per query
    unsigned int* lb2 = lb + kMmaMaxNQ;                       // [kMmaMaxNQ] ... and its second best
    float* tau_f = (float*)(lb2 + kMmaMaxNQ);                 // [kMmaMaxNQ] tau as a float (-inf = admit everything)
    float* lb_f = tau_f + kMmaMaxNQ;                          // [kMmaMaxNQ] lb2 as a float (may lag behind lb2: only a hint)
    unsigned int* rstage = (unsigned int*)(lb_f + kMmaMaxNQ); // [kMaxGrid] refresher scratch: one query's column of second-bests
    uint4* hitq = (uint4*)(rstage + kMaxGrid);                // [8][kMmaHitQueue] (query, row, score bits): staged hits
    int* hitn = (int*)(hitq + 8 * kMmaHitQueue);              // [8] entries staged per epilogue warp

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t n_tiles_all = (p.n_rows + kMmaRows - 1) / kMmaRows;
    const uint32_t n_tiles = p.mode == 0 ? (n_tiles_all < gridDim.x ? n_tiles_all : gridDim.x) : n_tiles_all;
    // tiles this CTA walks; a pair walks as many as its even CTA (the odd one may end on a padding tile)
    const uint32_t first_tile = PAIR ? (blockIdx.x & ~1u) : blockIdx.x;
    const uint32_t n_iters = first_tile < n_tiles ? (n_tiles - first_tile + gridDim.x - 1) / gridDim.x : 0u;
    const uint32_t tmem_cols = 2 * p.nqp <= 32 ? 32 : (2 * p.nqp <= 64 ? 64 : (2 * p.nqp <= 128 ? 128 : (2 * p.nqp <= 256 ? 256 : 512)));

    if (tid == 0) {
        for (int s = 0; s < kMmaStages; s++) {
            mbar_init(full + s, 1);
            mbar_init(conv + s, 4);        // one arrive per converter warp
            mbar_init(empty + s, 1);       // tcgen05.commit
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(acc_full + s, 1);    // tcgen05.commit
            mbar_init(acc_empty + s, (TERMS == 1 ? 8 : 4) * (PAIR ? 2 : 1));   // one arrive per epilogue warp (of both CTAs of a pair)
        }
        fence_mbar_init();
        flags[0] = 0;
        flags[1] = p.mode == 0 ? 1 : 0;    // the bound-only pass needs no thresholds
        for (int i = 0; i < 8; i++) hitn[i] = 0;
    }
    for (int i = tid; i < kMmaMaxNQ; i += kMmaThreads) {
        tau[i] = 0u;
        tau_f[i] = -INFINITY;
        // Both passes start their own tally: the full pass walks the bound-only pass's tile AGAIN, and a row
        // offered twice would count as two rows (its score would become the "second best" as well).  What the
        // bound-only pass found lives on in gbound[], which only rises.
        lb[i] = 0u;
        lb2[i] = 0u;
        lb_f[i] = -INFINITY;
    }
    if (warp == 1) { if constexpr (PAIR) tmem_alloc_pair(tmem_base_s, tmem_cols); else tmem_alloc(tmem_base_s, tmem_cols); }
    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();   // the peer's barriers are initialised too
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_s;

    if (warp == 0) {
        // ============================ TMA producer ============================
        // (whole warp on uniform values, one elected lane issues -- see the MMA issuer)
        if (elect_one()) {
            tma_prefetch_desc(&tmap_k);
            tma_prefetch_desc(&tmap_q);
        }
        __syncwarp();
        const uint64_t pol_stream = l2_policy_evict_first();
        const uint64_t pol_keep = l2_policy_evict_normal();
        const uint32_t full0 = PAIR ? mapa_u32(smem_u32(full), 0) : 0u;   // the leader's full[0], as a cluster address
        int stage = 0; uint32_t phase = 0;
        for (uint32_t it = 0; it < n_iters; it++) {
            const uint32_t tile = blockIdx.x + it * gridDim.x;
            for (int c = 0; c < p.n_chunks; c++) {
                mbar_wait(empty + stage, phase ^ 1u);
                if (elect_one()) {
                    if constexpr (PAIR) {
                        // the leader's barrier counts both CTAs' bytes; the peer only adds its transfers to it
                        if (crank == 0) mbar_arrive_expect_tx(full + stage, (uint32_t)(2 * (a_bytes + b_bytes)));
                        const uint32_t bar = full0 + (uint32_t)stage * 8u;
                        tma_load_2d_pair(a_hi(stage), &tmap_k, bar, c * bc, (int)(tile * kMmaRows), pol_stream);
                        tma_load_2d_pair(b_hi(stage), &tmap_q, bar, c * bc, (int)(crank * (uint32_t)(p.nqp / 2)), pol_keep);
                    } else {
                        mbar_arrive_expect_tx(full + stage, (uint32_t)(a_bytes + (TERMS == 3 ? 2 : 1) * b_bytes));
                        tma_load_2d(a_hi(stage), &tmap_k, full + stage, c * bc, (int)(tile * kMmaRows), pol_stream);
                        tma_load_2d(b_hi(stage), &tmap_q, full + stage, c * bc, 0, pol_keep);
                        if constexpr (TERMS == 3)
                            tma_load_2d(b_lo(stage), &tmap_q, full + stage, c * bc, p.nqp, pol_keep);
                    }
                }
                __syncwarp();
                if (++stage == kMmaStages) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        // The WHOLE warp walks the loop on warp-uniform values and one elected lane issues: descriptors then
        // live in uniform registers and an MMA costs a handful of instructions.  (Issuing from inside a
        // `lane == 0` branch made the compiler wrap every tcgen05 instruction in a uniformity loop: ~35
        // instructions per MMA, and the issuing thread -- not the tensor core -- set the kernel's pace.)
        if (crank == 0) {
            constexpr int kSteps = (TERMS == 3 ? 16 : 32) / 8;       // UMMA K = 8 TF32 = 32 bytes along the swizzled row
            const uint32_t idesc = umma_idesc_tf32(p.nqp, PAIR ? 256 : 128);
            // a stage's descriptors are stage 0's plus a multiple of the stage size: the address field counts 16-byte units
            const uint64_t a_desc0 = umma_desc(smem_u32(a_hi(0)), wide), b_desc0 = umma_desc(smem_u32(b_hi(0)), wide);
            const uint64_t stage_step = (uint64_t)(stage_bytes >> 4);
            const uint64_t a_lo_off = (uint64_t)(a_bytes >> 4), b_lo_off = (uint64_t)(b_bytes >> 4);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (uint32_t it = 0; it < n_iters; it++) {
                mbar_wait(acc_empty + acc, acc_phase ^ 1u);     // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.nqp);
                for (int c = 0; c < p.n_chunks; c++) {
                    mbar_wait(full + stage, phase);             // key and query boxes landed
                    if constexpr (TERMS == 3) mbar_wait(conv + stage, phase);   // hi / lo key tiles written and fenced
                    tc_fence_after();
                    if (elect_one()) {
                        const uint64_t ad = a_desc0 + (uint64_t)stage * stage_step, bd = b_desc0 + (uint64_t)stage * stage_step;
                        if (!(p.debug & 2)) {
#pragma unroll
                            for (int k = 0; k < kSteps; k++) {
                                const uint64_t off = (uint64_t)(2 * k);     // 32 bytes
                                if constexpr (PAIR) umma_tf32_pair(d_tmem, ad + off, bd + off, idesc, (c | k) != 0);
                                else umma_tf32(d_tmem, ad + off, bd + off, idesc, (c | k) != 0);
                                if constexpr (TERMS == 3) {
                                    umma_tf32(d_tmem, ad + off, bd + b_lo_off + off, idesc, 1u);
                                    umma_tf32(d_tmem, ad + a_lo_off + off, bd + off, idesc, 1u);
                                }
                            }
                        }
                        if constexpr (PAIR) umma_commit_pair(empty + stage); else umma_commit(empty + stage);   // stage is free once these MMAs retire
                    }
                    __syncwarp();
                    if (++stage == kMmaStages) { stage = 0; phase ^= 1u; }
                }
                if (elect_one()) {
                    if constexpr (PAIR) umma_commit_pair(acc_full + acc); else umma_commit(acc_full + acc);     // accumulator complete
                }
                __syncwarp();
                if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
            }
        }
    } else if (warp == 2) {
        // ============================ bound refresher ============================
        // Every CTA tracks its two best scores per query.  The r-th largest of the CTAs' SECOND bests is a
        // lower bound of the (2r)-th best score of the store (r CTAs hold two rows each at least that good),
        // and 2r >= need (k for the single-TF32 filter, whose error band hangs below the k-th best; k + margin
        // for 3xTF32).  Far tighter than the minimum of the CTAs' bests, which the unluckiest CTA sets.
        // The work is spread: CTA b computes the bound of queries b, b + grid, ... into gbound[] (it only
        // rises: atomicMax), and every CTA refreshes its thresholds from gbound[] -- 256 loads, not 148 x 256.
        if (p.mode == 1) {
            for (;;) {
                for (int q = lane; q < p.nq; q += 32) p.lbest2[(size_t)blockIdx.x * kMmaMaxNQ + q] = lb2[q];
                for (int q = (int)blockIdx.x; q < p.nq; q += (int)gridDim.x) {
                    const unsigned int v = mma_rth_largest(p.lbest2 + q, (int)gridDim.x, p.r_need, lane, rstage);
                    if (lane == 0 && v != 0u) atomicMax(p.gbound + q, v);
                }
                for (int q = lane; q < p.nq; q += 32) {
                    unsigned int m = ((volatile unsigned int*)p.gbound)[q];
                    if (m == 0u) continue;                       // no bound yet
                    // admit everything within 2 eps below the bound (eps = 0 when the filter is 3xTF32)
                    const float e2 = p.eps ? 2.0f * p.eps[q] : 0.0f;
                    if (e2 > 0.0f) m = score_to_order(order_to_score(m) - e2);
                    if (m > tau[q]) {
                        ((volatile unsigned int*)tau)[q] = m;
                        ((volatile float*)tau_f)[q] = order_to_score(m);
                    }
                }
                __syncwarp();
                if (lane == 0) flags[1] = 1;
                int d = (lane == 0) ? flags[0] : 0;
                if (__shfl_sync(0xFFFFFFFFu, d, 0)) break;
                __nanosleep((unsigned)p.refresh_ns);
            }
        }
    } else if (warp >= 4 && (warp < 8 || TERMS == 1)) {
        // ============================ epilogue ============================
        // Warps 4-7, and with the single-TF32 filter warps 8-11 too (they have nothing to convert): warp w may
        // read TMEM lanes 32 (w % 4) .. +31, so the two groups take alternate 32-query column blocks of the same rows.
        constexpr int kEpiWarps = TERMS == 1 ? 8 : 4;
        const int wq = warp & 3;                                // TMEM lane quadrant this warp may read
        const int ew = warp - 4;                                // epilogue warp index: its hit queue
        const int cgrp = ew >> 2;                               // which alternate column blocks
        const int row_in_tile = wq * 32 + lane;
        // One element cleared a threshold (rare once the bounds have tightened): update this CTA's best for the
        // query and, on the full pass, stage the candidate.  Runs only for rows whose mask bit is set, after the
        // branch-free pass over the block; all its state is addressed as shared memory (no generic atomics).
        int* const my_hitn = hitn + ew;
        uint4* const my_hitq = hitq + ew * kMmaHitQueue;
        auto note_best = [&](int q, unsigned int o) {
            if (o > ((volatile unsigned int*)lb2)[q]) {
                // keep the CTA's two best: the new best goes to lb, whichever of (old best, this score) is not
                // the best any more is offered to lb2.  Every score but the largest ends up offered to lb2.
                const unsigned int old1 = atomicMax(lb + q, o);
                const unsigned int lower = old1 < o ? old1 : o;
                const unsigned int old2 = atomicMax(lb2 + q, lower);
                ((volatile float*)lb_f)[q] = order_to_score(old2 > lower ? old2 : lower);
            }
        };
        auto on_hit = [&](int q, uint32_t row, float sc, bool note) {
            if (q >= p.nq) return;
            const unsigned int o = score_to_order(sc);
            if (note) note_best(q, o);
            if (p.mode == 1 && o >= ((volatile unsigned int*)tau)[q]) {
                // A global atomic per hit would stall the warp ~1 us each time (its result is needed for the
                // store): stage the hit, the warp flushes all of them after the tile.
                const int hp = atomicAdd(my_hitn, 1);
                if (hp < kMmaHitQueue) {
                    my_hitq[hp] = make_uint4((unsigned)q, row, __float_as_uint(sc), 0u);
                } else {   // queue full (thresholds still loose): the slow path, rare
                    const unsigned int pos = atomicAdd(p.cand_cnt + q, 1u);
                    if (pos < (unsigned)kMmaCandCap) p.cand[(size_t)q * kMmaCandCap + pos] = pack_cand(sc, row);
                }
            }
        };
        {   // first thresholds are in place (bounded wait: a protocol bug must trap, not hang)
            const long long t0 = clock64();
            while (flags[1] == 0) {
                __nanosleep(200);
                if (clock64() - t0 > 4000000000ll) { printf("tfp: mma epilogue never saw thresholds\n"); __trap(); }
            }
        }
        int acc = 0; uint32_t acc_phase = 0;
        for (uint32_t it = 0; it < n_iters; it++) {
            const uint32_t tile = blockIdx.x + it * gridDim.x;
            const uint32_t row = tile * kMmaRows + (uint32_t)row_in_tile;
            const bool valid = row < p.n_rows;
            const float inv = (p.f64 && valid) ? (float)(1.0 / __ldg(p.knorm + row)) : 1.0f;
            mbar_wait(acc_full + acc, acc_phase);
            tc_fence_after();
            for (int col0 = cgrp * 32; col0 < p.nqp; col0 += 32 * (kEpiWarps / 4)) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(acc * p.nqp + col0), v);
                // Thresholds and bests of these 32 queries as floats: 16 vector loads up front, then two float
                // compares per element folded into one bit of a mask -- no branch per element.  Only a row with
                // a bit set (rare once the bounds are tight) walks its bits; that path works on the exact
                // orderable words.  A stale float is harmless: a lower threshold or best only sets a few more bits.
                float4 tq[8], bq[8];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    tq[i] = __uint4_as_float4(lds_volatile_v4((const unsigned int*)tau_f + col0 + 4 * i));
                    bq[i] = __uint4_as_float4(lds_volatile_v4((const unsigned int*)lb_f + col0 + 4 * i));
                }
                uint32_t mask = 0u;
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const float tj = j % 4 == 0 ? tq[j / 4].x : (j % 4 == 1 ? tq[j / 4].y : (j % 4 == 2 ? tq[j / 4].z : tq[j / 4].w));
                    const float bj = j % 4 == 0 ? bq[j / 4].x : (j % 4 == 1 ? bq[j / 4].y : (j % 4 == 2 ? bq[j / 4].z : bq[j / 4].w));
                    const float sj = __uint_as_float(v[j]) * inv;
                    mask |= ((sj >= tj) | (sj > bj)) ? (1u << j) : 0u;      // both false for NaN
                }
                if (!valid || (p.debug & 1)) mask = 0u;
                // The first tile of a pass meets empty bests: every element is a "hit", and each costs a chain of
                // dependent shared-memory atomics on its query's two bests (ncu: 60 % of the bound-only pass's samples
                // sat in that code, another 26 % in warps waiting for it at the end).  While a block is that dense
                // the warp finds each query's two best of its 32 rows with two redux steps -- independent across the
                // 32 queries, so they pipeline -- lane j keeps those of query col0 + j, and ONE round of atomics, 32
                // queries side by side, offers them (the CTA's two best are among the warps' two best).
                if (__any_sync(0xFFFFFFFFu, __popc(mask) >= 12)) {
                    unsigned int my_m1 = 0u, my_m2 = 0u;
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const float sj = __uint_as_float(v[j]) * inv;
                        const unsigned int o = (((mask >> j) & 1u) && col0 + j < p.nq) ? score_to_order(sj) : 0u;
                        const unsigned int m1 = __reduce_max_sync(0xFFFFFFFFu, o);
                        const int first = __ffs(__ballot_sync(0xFFFFFFFFu, o == m1)) - 1;
                        const unsigned int m2 = __reduce_max_sync(0xFFFFFFFFu, lane == first ? 0u : o);
                        if (lane == j) { my_m1 = m1; my_m2 = m2; }
                        if (p.mode == 1 && o != 0u) on_hit(col0 + j, row, sj, false);
                    }
                    if (my_m1 != 0u) note_best(col0 + lane, my_m1);
                    if (my_m2 != 0u) note_best(col0 + lane, my_m2);
                } else if (mask) {
#pragma unroll
                    for (int j = 0; j < 32; j++)
                        if (mask & (1u << j)) on_hit(col0 + j, row, __uint_as_float(v[j]) * inv, true);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {                                    // the tensor core may refill this accumulator now
                if (PAIR && crank != 0) mbar_arrive_cluster(mapa_u32(smem_u32(acc_empty + acc), 0));
                else mbar_arrive(acc_empty + acc);
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
            // flush the staged hits: 32 global atomics in flight at a time instead of one
            {
                int nh = *((volatile int*)(hitn + ew));
                nh = nh < kMmaHitQueue ? nh : kMmaHitQueue;
                for (int i = lane; i < nh; i += 32) {
                    const uint4 e = hitq[ew * kMmaHitQueue + i];
                    const unsigned int pos = atomicAdd(p.cand_cnt + e.x, 1u);
                    if (pos < (unsigned)kMmaCandCap) p.cand[(size_t)e.x * kMmaCandCap + pos] = pack_cand(__uint_as_float(e.z), e.y);
                }
                __syncwarp();
                if (lane == 0) hitn[ew] = 0;
                __syncwarp();
            }
        }
        // publish this CTA's bests (the bound-only pass exists to produce exactly this)
        named_bar_sync(2, kEpiWarps * 32);
        for (int q = tid - 128; q < p.nq; q += kEpiWarps * 32) p.lbest2[(size_t)blockIdx.x * kMmaMaxNQ + q] = lb2[q];
        if (tid == 128) flags[0] = 1;
    } else if (TERMS == 3 && warp >= 8) {
        // ============================ converters (3xTF32 only) ============================
        const int ct = tid - 256;                               // 0..127
        int stage = 0; uint32_t phase = 0;
        for (uint32_t it = 0; it < n_iters; it++) {
            for (int c = 0; c < p.n_chunks; c++) {
                mbar_wait(full + stage, phase);
                float4* hi4 = reinterpret_cast<float4*>(a_hi(stage));
                float4* lo4 = reinterpret_cast<float4*>(a_lo(stage));
#pragma unroll
                for (int i = 0; i < (int)(a_bytes / 16 / 128); i++) {   // element-wise: the swizzle does not matter
                    float4 a = hi4[ct + i * 128], h, l;
                    tf32_split(a.x, h.x, l.x);
                    tf32_split(a.y, h.y, l.y);
                    tf32_split(a.z, h.z, l.z);
                    tf32_split(a.w, h.w, l.w);
                    hi4[ct + i * 128] = h;
                    lo4[ct + i * 128] = l;
                }
                fence_proxy_async_smem();                        // generic-proxy writes -> visible to the tensor core
                __syncwarp();
                if (lane == 0) mbar_arrive(conv + stage);
                if (++stage == kMmaStages) { stage = 0; phase ^= 1u; }
            }
        }
    }
    // ---- teardown: every MMA has retired (the epilogue consumed the last accumulator) ----
    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();   // neither CTA of a pair leaves while the other may touch it
    tc_fence_after();
    if (warp == 1) { if constexpr (PAIR) tmem_dealloc_pair(tmem_base, tmem_cols); else tmem_dealloc(tmem_base, tmem_cols); }
}

// ---- finalize: per query, best k+margin candidates by approximate score, exact re-score, top-k ----
struct FinalizeParams {
    const float* keys; const int64_t* ids; const double* knorm;
    uint32_t n_rows; int dim;
    const float* q; const QInfo* qinfo; int q_first;     // original (unsplit) staged queries; q_first: the staged query of output slot out_first
    const int* qmap;                                     // [nq] staged query served by each query of this pass
    const uint64_t* cand; const unsigned int* cand_cnt;
    int topk;
    int out_first;
    uint64_t* out_cand; int64_t* out_ids; float* out_sims; uint32_t* out_slots; int* out_n;
    uint8_t* out_rec; int rec_bytes; int rec_topk;
    int* status;                                         // [nq] 0 ok, else the host re-runs the query exactly
    const float* eps;                                    // [nq] error bound of the approximate scores (TERMS = 1)
};

__device__ __forceinline__ void block_sort_desc(uint64_t* buf, int P, int tid, int nthreads) {
    for (int k2 = 2; k2 <= P; k2 <<= 1) {
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (P >> 1); i += nthreads) {
                int lo = ((i / j) * (j << 1)) + (i % j), hi = lo + j;
                uint64_t x = buf[lo], y = buf[hi];
                bool desc = ((lo & k2) == 0);
                if ((x < y) == desc && x != y) { buf[lo] = y; buf[hi] = x; }
            }
            __syncthreads();
        }
    }
}

// The rows to re-score are found by SELECTION, not by sorting the whole list: the 32-bit score words are
// staged in shared memory once (64 KiB: two blocks per SM, so the 256 queries of a batch are resident at
// once), a 4-pass radix select over them gives the approximate score of the rank that matters, and one more
// pass collects every candidate that reaches the cut (only those are re-read from the list).
constexpr int kFinThreads = 256;       // few rows are re-scored (tens to hundreds); 128 registers each for their pipelined row walk
constexpr int kFinWarps = kFinThreads / 32;
constexpr size_t kFinDynSmem = (size_t)kMmaCandCap * 4;
template <bool F64, int TERMS>
__global__ void __launch_bounds__(kFinThreads, 2)
mma_finalize_kernel(const FinalizeParams p) {
    extern __shared__ unsigned int skey[];               // [n] score words of the candidate list
    __shared__ uint64_t fbuf[kMmaRescoreMax];            // approximate words of the rows to re-score
    __shared__ uint64_t exact[kMmaRescoreMax];           // exact words of the re-scored rows
    __shared__ unsigned int hist[kFinWarps][256];        // one histogram per warp: no cross-warp contention on hot bins
    __shared__ unsigned int s_prefix, s_rank, s_cnt;
    const int q = blockIdx.x, tid = threadIdx.x;
    const unsigned int cnt = p.cand_cnt[q];
    const int n = cnt < (unsigned)kMmaCandCap ? (int)cnt : kMmaCandCap;
    const uint64_t* cand = p.cand + (size_t)q * kMmaCandCap;
    const int k = p.topk;
    const int lim = (uint32_t)k < p.n_rows ? k : (int)p.n_rows;
    // the rank whose approximate score decides the cut: k-th best (single TF32: the error band hangs below
    // it), or (k + margin)-th best (3xTF32: exactly that many are re-scored, ties included)
    const int want = TERMS == 3 ? (n < k + kMmaMargin ? n : k + kMmaMargin) : (n < lim ? n : lim);
    if (tid == 0) { s_prefix = 0u; s_rank = (unsigned)want; s_cnt = 0u; }
#pragma unroll 8
    for (int i = tid; i < n; i += kFinThreads) skey[i] = (unsigned int)(__ldcg(cand + i) >> 32);   // independent loads: many in flight
    __syncthreads();
    if (want >= 1) {
        unsigned int* myhist = hist[tid >> 5];
        for (int shift = 24; shift >= 0; shift -= 8) {
            for (int i = tid; i < kFinWarps * 256; i += kFinThreads) (&hist[0][0])[i] = 0u;
            __syncthreads();
            const unsigned int prefix = s_prefix;
            for (int i = tid; i < n; i += kFinThreads) {
                const unsigned int key = skey[i];
                if (shift == 24 || ((key ^ prefix) >> (shift + 8)) == 0u) atomicAdd(&myhist[(key >> shift) & 255u], 1u);
            }
            __syncthreads();
            for (int b = tid; b < 256; b += kFinThreads) {   // fold the per-warp histograms into hist[0]
                unsigned int sum = 0;
                for (int w = 0; w < kFinWarps; w++) sum += hist[w][b];
                hist[0][b] = sum;
            }
            __syncthreads();
            if (tid < 32) {   // warp 0: lane l owns bins 255 - 8l .. 248 - 8l (descending keys)
                unsigned int local = 0;
                for (int j = 0; j < 8; j++) local += hist[0][255 - 8 * tid - j];
                unsigned int incl = local;
                for (int d = 1; d < 32; d <<= 1) {
                    const unsigned int v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                    if (tid >= d) incl += v;
                }
                const unsigned int rank = s_rank;
                const bool mine = incl >= rank && incl - local < rank;   // the rank-th largest falls into my bins
                if (mine) {
                    unsigned int acc = incl - local;
                    for (int j = 0; j < 8; j++) {
                        const int b = 255 - 8 * tid - j;
                        if (acc + hist[0][b] >= rank) { s_prefix = prefix | ((unsigned int)b << shift); s_rank = rank - acc; break; }
                        acc += hist[0][b];
                    }
                }
            }
            __syncthreads();
        }
    }
    // every candidate whose score word reaches `cut` is re-scored
    unsigned int cut = 0u;                               // n <= lim: take them all
    if (TERMS == 3) cut = want >= 1 ? s_prefix : 0u;
    else if (n > lim) {
        // every row whose approximate score is within 2 eps of the k-th best approximate score:
        // a row below that band is below the k-th exact score (as long as eps bounds the filter's error)
        cut = score_to_order(order_to_score(s_prefix) - 2.0f * p.eps[q]);
    }
    for (int i = tid; i < n; i += kFinThreads) {
        if (skey[i] >= cut) {
            const unsigned int pos = atomicAdd(&s_cnt, 1u);
            if (pos < (unsigned)kMmaRescoreMax) fbuf[pos] = __ldcg(cand + i);
        }
    }
    __syncthreads();
    const int C_all = (int)s_cnt;
    const int C = C_all < kMmaRescoreMax ? C_all : kMmaRescoreMax;
    const uint32_t cut_order = cut;                      // 3xTF32: rows not re-scored all score below this
    // exact re-score, one thread per candidate row, the reference's arithmetic (scan.cuh Acc<>)
    const int sq = p.qmap[q];                            // the staged query this block serves
    const float* qv = p.q + (size_t)sq * p.dim;
    const double mag1 = F64 ? p.qinfo[sq].mag1 : 0.0;
    // The query goes to shared memory (the score words there are dead now): every thread reads all of it, and
    // left in global memory its lines are pushed out of L1 by the key rows streaming through -- each chunk of
    // the walk then waited a full L2 round trip for its query values (measured: 78 % of this kernel).
    const bool q_in_smem = (size_t)p.dim * 4 <= kFinDynSmem;
    float4* qs4 = reinterpret_cast<float4*>(skey);
    if (q_in_smem)
        for (int i = tid; i < (p.dim >> 2); i += kFinThreads) qs4[i] = __ldg(reinterpret_cast<const float4*>(qv) + i);
    __syncthreads();
    // Every thread walks its own (cold) row, one add after the other: what costs is memory latency, not
    // arithmetic.  The walk is software-pipelined in 128-byte chunks -- the next chunk's loads are in flight
    // while this chunk's 32 adds run.  (What remains is HBM's rate for ~190 scattered 4 KiB rows per query:
    // the same walk over adjacent rows runs 0.14 ms faster per 256-query batch; an L2 prefetch of the whole
    // row up front was measured and made it slightly slower.)
    // dim % 4 == 0 on this path; the adds stay in index order.  (A lambda, so that the query pointer keeps
    // its address space: shared-memory loads when the query is staged there.)
    auto rescore = [&](uint32_t slot, auto qv4) -> uint64_t {
        const float4* kr4 = reinterpret_cast<const float4*>(p.keys + (size_t)slot * p.dim);
        const int n4 = p.dim >> 2;
        constexpr int CH = 8;
        const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float4 cur[CH], nxt[CH];
#pragma unroll
        for (int j = 0; j < CH; j++) cur[j] = j < n4 ? __ldcs(kr4 + j) : zero4;
        Acc<F64> acc;
        acc.zero();
        for (int base = 0; base < n4; base += CH) {
#pragma unroll
            for (int j = 0; j < CH; j++) nxt[j] = base + CH + j < n4 ? __ldcs(kr4 + base + CH + j) : zero4;
#pragma unroll
            for (int j = 0; j < CH; j++) {
                if (base + j < n4) {
                    const float4 qq = qv4[base + j];
                    acc.add(qq.x, cur[j].x); acc.add(qq.y, cur[j].y); acc.add(qq.z, cur[j].z); acc.add(qq.w, cur[j].w);
                }
            }
#pragma unroll
            for (int j = 0; j < CH; j++) cur[j] = nxt[j];
        }
        return pack_cand(finish_score<F64>(acc, mag1, F64 ? p.knorm[slot] : 1.0), slot);
    };
    // The walk as written above has lane l of a warp read 16 bytes of ITS row per load: 32 different lines per
    // instruction, 256 L1 wavefronts per warp for every 128-byte step of the 32 rows (ncu, round 2: 83 % of this
    // kernel's samples sit in the walk and the barrier behind it, at 1.3 TB/s).  A warp therefore fetches the step
    // of its 32 rows TOGETHER -- 8 lanes per row, 4 whole 128-byte lines per instruction, 32 wavefronts per step --
    // and hands every lane its own row's 128 bytes through a padded shared-memory tile (144-byte rows: both the
    // 128-bit stores and the 128-bit loads are conflict-free).  Each lane still adds ITS row in index order: the
    // arithmetic is unchanged.  The tile lives behind the staged query in the (dead) score-word array.
    constexpr int kStgRow = 9;                           // float4 per staged row: 8 of data + 1 of padding
    const size_t stg_off = (((size_t)p.dim * 4 + 15) & ~(size_t)15);
    const bool coop = q_in_smem && stg_off + (size_t)kFinWarps * 32 * kStgRow * 16 <= kFinDynSmem;
    if (coop) {
        const int lane = tid & 31, wrp = tid >> 5;
        const int lr = lane >> 3, lc = lane & 7;
        float4* stg = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(skey) + stg_off) + (size_t)wrp * 32 * kStgRow;
        const float4* k4 = reinterpret_cast<const float4*>(p.keys);
        const float4* qv4 = (const float4*)qs4;
        const int n4 = p.dim >> 2;
        const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (int t0 = wrp * 32; t0 < kMmaRescoreMax; t0 += kFinThreads) {
            const int t = t0 + lane;
            if (t0 >= C) { exact[t] = 0ull; continue; }  // (warp-uniform)
            const bool valid = t < C;
            const uint32_t my_slot = valid ? cand_slot(fbuf[t]) : 0u;
            const unsigned vmask = __ballot_sync(0xFFFFFFFFu, valid);
            uint32_t rslot[8];
#pragma unroll
            for (int i = 0; i < 8; i++) rslot[i] = __shfl_sync(0xFFFFFFFFu, my_slot, i * 4 + lr);
            float4 nxt[8], cur[8];
            auto fetch = [&](int base) {
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const bool on = ((vmask >> (i * 4 + lr)) & 1u) && base + lc < n4;
                    nxt[i] = on ? __ldcs(k4 + (size_t)rslot[i] * n4 + base + lc) : zero4;
                }
            };
            fetch(0);
            Acc<F64> acc;
            acc.zero();
            for (int base = 0; base < n4; base += 8) {
#pragma unroll
                for (int i = 0; i < 8; i++) stg[(i * 4 + lr) * kStgRow + lc] = nxt[i];
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; j++) cur[j] = stg[lane * kStgRow + j];
                __syncwarp();
                if (base + 8 < n4) fetch(base + 8);      // in flight while this step's 32 adds run
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (base + j < n4) {
                        const float4 qq = qv4[base + j];
                        acc.add(qq.x, cur[j].x); acc.add(qq.y, cur[j].y); acc.add(qq.z, cur[j].z); acc.add(qq.w, cur[j].w);
                    }
                }
            }
            exact[t] = valid ? pack_cand(finish_score<F64>(acc, mag1, F64 ? p.knorm[my_slot] : 1.0), my_slot) : 0ull;
        }
    } else {
    for (int t = tid; t < kMmaRescoreMax; t += blockDim.x) {
        uint64_t w = 0ull;
        if (t < C) {
            const uint32_t slot = cand_slot(fbuf[t]);
            if (q_in_smem) w = rescore(slot, (const float4*)qs4);
            else w = rescore(slot, reinterpret_cast<const float4*>(qv));
        }
        exact[t] = w;
    }
    }
    __syncthreads();
    int PE = 2;
    while (PE < C) PE <<= 1;
    __syncthreads();
    block_sort_desc(exact, PE, tid, blockDim.x);
    const int m = C < lim ? C : lim;
    const int o = p.out_first + (sq - p.q_first);
    for (int i = tid; i < k; i += blockDim.x) {
        size_t at = (size_t)o * k + i;
        if (i < m) {
            const uint64_t c = exact[i];
            const uint32_t slot = cand_slot(c);
            const int64_t id = p.ids[slot];
            p.out_cand[at] = c; p.out_sims[at] = cand_score(c); p.out_slots[at] = slot; p.out_ids[at] = id;
            record_write(p.out_rec + (size_t)o * p.rec_bytes, p.rec_topk, i, id, cand_score(c));
        } else {
            p.out_cand[at] = 0ull; p.out_sims[at] = 0.0f; p.out_slots[at] = 0xFFFFFFFFu; p.out_ids[at] = -1;
        }
    }
    if (tid == 0) {
        p.out_n[o] = m;
        *(int*)(p.out_rec + (size_t)o * p.rec_bytes) = m;
        int st = 0;
        if (cnt > (unsigned)kMmaCandCap) st = 1;                     // candidate list overflowed
        else if (m < lim) st = 2;                                    // came out short
        else if (TERMS == 1 && C_all > kMmaRescoreMax) st = 4;       // too many rows inside the error band
        else if (TERMS == 3 && (uint32_t)C < p.n_rows && m > 0) {
            // Rows that were NOT re-scored all have an approximate score <= `cut` (the last re-scored
            // candidate; rows never listed are below the threshold, which is below it too).  They are
            // irrelevant iff cut + (3xTF32 error) is still below the k-th exact score.
            const float kth = cand_score(exact[m - 1]);
            const float nxt = order_to_score(cut_order);
            if (!(nxt + 2e-5f * (fabsf(kth) + 1.0f) < kth)) st = 3;  // too close to call: exact re-run
        }
        p.status[q] = st;
    }
}

}  // namespace tfp
