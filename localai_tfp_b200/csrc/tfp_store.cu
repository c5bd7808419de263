// tfp_store.cu -- host side of libtfpstore.so and the C ABI of include/tfp_store.h.
//
// Mirrors the reference's Store (backend/go/stores/store.go:17-30): the sticky
// normalised flag, the latched key length, the error texts, the set / delete /
// get / find semantics -- with the keys resident in B200 HBM and every loop over
// keys running as a CUDA kernel (scan.cuh, select.cuh, ingest.cuh).
// There is no CPU fallback: if CUDA is unusable every entry point fails.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <fcntl.h>
#include <unistd.h>
#include <memory>
#include <mutex>
#include <pthread.h>
#include <shared_mutex>
#include <string>
#include <unordered_set>
#include <vector>

#include "../../include/tfp_store.h"
#include "host_index.hpp"
#include "ingest.cuh"
#include "scan.cuh"
#include "mma_scan.cuh"

using namespace tfp;

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CU_OK(expr)                                                                                       \
    do {                                                                                                  \
        cudaError_t e__ = (expr);                                                                         \
        if (e__ != cudaSuccess)                                                                           \
            return fail(-1, "CUDA error %s at %s:%d: %s", cudaGetErrorName(e__), __FILE__, __LINE__, #expr); \
    } while (0)

extern "C" const char* tfp_last_error(void) { return g_err.c_str(); }
extern "C" const char* tfp_version(void) { return "libtfpstore 0.1 (sm_100a; TMA scan, fused top-k)"; }

// ---------------------------------------------------------------------------
// Reader-writer lock that lets a waiting writer in ahead of new readers.  Finds take the store's lock shared, one
// after the other and from many threads at once; std::shared_mutex (glibc's default rwlock) prefers readers, and a
// StoresSet that only needs the lock for an instant to publish its rows would wait for as long as finds keep coming.
// Non-recursive: a thread never takes the same lock shared twice.
// ---------------------------------------------------------------------------
class RwMutex {
    pthread_rwlock_t l_;
public:
    RwMutex() {
        pthread_rwlockattr_t a;
        pthread_rwlockattr_init(&a);
        pthread_rwlockattr_setkind_np(&a, PTHREAD_RWLOCK_PREFER_WRITER_NONRECURSIVE_NP);
        pthread_rwlock_init(&l_, &a);
        pthread_rwlockattr_destroy(&a);
    }
    ~RwMutex() { pthread_rwlock_destroy(&l_); }
    RwMutex(const RwMutex&) = delete;
    RwMutex& operator=(const RwMutex&) = delete;
    void lock() { pthread_rwlock_wrlock(&l_); }
    bool try_lock() { return pthread_rwlock_trywrlock(&l_) == 0; }
    void unlock() { pthread_rwlock_unlock(&l_); }
    void lock_shared() { pthread_rwlock_rdlock(&l_); }
    bool try_lock_shared() { return pthread_rwlock_tryrdlock(&l_) == 0; }
    void unlock_shared() { pthread_rwlock_unlock(&l_); }
};

// ---------------------------------------------------------------------------
// small RAII-less buffers (explicit release; the store owns them)
// ---------------------------------------------------------------------------
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    int reserve(size_t want) {
        if (want <= bytes) return 0;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        size_t cap = std::max<size_t>(want, 256);
        CU_OK(cudaMalloc(&p, cap));
        bytes = cap;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    template <class T> T* as() const { return (T*)p; }
};
struct PinBuf {
    void* p = nullptr;
    size_t bytes = 0;
    int reserve(size_t want) {
        if (want <= bytes) return 0;
        if (p) cudaFreeHost(p);
        p = nullptr; bytes = 0;
        size_t cap = std::max<size_t>(want, 4096);
        CU_OK(cudaMallocHost(&p, cap));
        bytes = cap;
        return 0;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
    template <class T> T* as() const { return (T*)p; }
};

// ---------------------------------------------------------------------------
// TMA descriptor encoding through the driver entry point (no -lcuda link, so
// the library still loads on a machine without a driver; it just cannot run)
// ---------------------------------------------------------------------------
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_tmapEncodeTiled g_encode = nullptr;
static std::once_flag g_encode_once;

static int get_encoder() {
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            g_encode = (PFN_tmapEncodeTiled)fn;
    });
    return g_encode ? 0 : fail(-2, "cuTensorMapEncodeTiled is not available from the CUDA driver");
}

// ---------------------------------------------------------------------------
// development knobs (environment; defaults are what ships)
// ---------------------------------------------------------------------------
struct Tuning {
    int refresh_ns = 4000;
    bool disable_lbest = false;
    int mma_min_queries = 3;       // batches at least this large go to the tensor-core path (measured break-even: 531 us per batch vs 896 us for 2 + 1 exact scans on 1M x 768)
    int64_t mma_min_rows = 65536;
    int mma_terms = 1;             // 1 = single-TF32 filter + re-score of its error band (3xTF32 as the second pass), 3 = 3xTF32 split first
    bool mma_pair = true;          // single-TF32 filter on CTA pairs (cta_group::2: half the query bytes per SM)
    int mma_serial = -1;           // tensor-core passes of different contexts are ordered one behind the other: 1 always, 0 never,
                                   // -1 when the keys are 2 GiB or more (see mma_pass_launch)
    bool mma_defer = true;         // tfp_find_ctx_launch leaves a tensor-core batch's status words to the context's next call
    int mma_debug = 0;             // development timing only (results wrong): 1 no epilogue hits, 2 no MMAs
    bool mma_trace = false;        // print the tensor-core launch shape
    bool scan_times = false;       // per-CTA time stamps of every exact scan, printed by tfp_find_ctx_sync
    int scan_stages = 0;           // cap on the exact scan's TMA ring depth (0 = as many as fit): fewer stages let two CTAs share an SM
    bool set_trace = false;        // print where StoresSet spends its time when the store closes
    bool xchg_stream_wait = false; // peer exchange: wait for the peers' words with stream memory operations instead of inside the merge kernel
};
static const Tuning& tuning() {
    static Tuning t = [] {
        Tuning x;
        if (const char* e = getenv("TFP_REFRESH_NS")) x.refresh_ns = atoi(e);
        if (const char* e = getenv("TFP_DISABLE_LBEST")) x.disable_lbest = atoi(e) != 0;
        if (const char* e = getenv("TFP_MMA_MIN_QUERIES")) x.mma_min_queries = atoi(e);
        if (const char* e = getenv("TFP_MMA_MIN_ROWS")) x.mma_min_rows = atoll(e);
        if (const char* e = getenv("TFP_MMA_TERMS")) x.mma_terms = atoi(e) == 3 ? 3 : 1;
        if (const char* e = getenv("TFP_MMA_PAIR")) x.mma_pair = atoi(e) != 0;
        if (const char* e = getenv("TFP_MMA_DEFER")) x.mma_defer = atoi(e) != 0;
        if (const char* e = getenv("TFP_MMA_SERIAL")) x.mma_serial = atoi(e);
        if (const char* e = getenv("TFP_MMA_DEBUG")) x.mma_debug = atoi(e);
        x.mma_trace = getenv("TFP_MMA_TRACE") != nullptr;
        x.scan_times = getenv("TFP_SCAN_TIMES") != nullptr;
        if (const char* e = getenv("TFP_SCAN_STAGES")) x.scan_stages = atoi(e);
        x.set_trace = getenv("TFP_SET_TRACE") != nullptr;
        if (const char* e = getenv("TFP_XCHG_STREAM_WAIT")) x.xchg_stream_wait = atoi(e) != 0;
        return x;
    }();
    return t;
}

// NCCL, resolved at run time (see tfp_find_ctx_launch_sharded); declared early because a find
// context owns a communicator.
struct NcclId { char internal[128]; };   // ncclUniqueId
struct NcclApi {
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};
static NcclApi g_nccl;
static std::once_flag g_nccl_once;

// ---------------------------------------------------------------------------
// Growable HBM arrays on CUDA virtual memory management: a large virtual range is reserved once,
// physical memory is mapped behind it as the store grows.  The base address never changes, no
// row is ever copied to grow, and the peak footprint is the data itself -- what lets one shard
// approach the 180 GB of a B200 (a realloc-and-copy array would need twice its size to grow).
// Driver entry points are resolved at run time (no -lcuda link).
// ---------------------------------------------------------------------------
struct VmmApi {
    CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*GetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    bool ok = false;
};
static VmmApi g_vmm;
static std::once_flag g_vmm_once;
static bool vmm_api() {
    std::call_once(g_vmm_once, [] {
        auto get = [](const char* name) -> void* {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
            return fn;
        };
        g_vmm.AddressReserve = (decltype(g_vmm.AddressReserve))get("cuMemAddressReserve");
        g_vmm.AddressFree = (decltype(g_vmm.AddressFree))get("cuMemAddressFree");
        g_vmm.Create = (decltype(g_vmm.Create))get("cuMemCreate");
        g_vmm.Release = (decltype(g_vmm.Release))get("cuMemRelease");
        g_vmm.Map = (decltype(g_vmm.Map))get("cuMemMap");
        g_vmm.Unmap = (decltype(g_vmm.Unmap))get("cuMemUnmap");
        g_vmm.SetAccess = (decltype(g_vmm.SetAccess))get("cuMemSetAccess");
        g_vmm.GetGranularity = (decltype(g_vmm.GetGranularity))get("cuMemGetAllocationGranularity");
        g_vmm.ok = g_vmm.AddressReserve && g_vmm.AddressFree && g_vmm.Create && g_vmm.Release && g_vmm.Map &&
                   g_vmm.Unmap && g_vmm.SetAccess && g_vmm.GetGranularity && getenv("TFP_DISABLE_VMM") == nullptr;
    });
    return g_vmm.ok;
}

struct VmmArray {
    CUdeviceptr base = 0;
    size_t reserved = 0, mapped = 0, gran = 0;
    int device = 0;
    std::vector<std::pair<CUmemGenericAllocationHandle, size_t>> chunks;

    bool active() const { return base != 0; }
    // Reserve `bytes` of address space (no memory yet).
    int reserve(int dev, size_t bytes) {
        device = dev;
        CUmemAllocationProp prop;
        memset(&prop, 0, sizeof(prop));
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = dev;
        if (g_vmm.GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0)
            return fail(-6, "cuMemGetAllocationGranularity failed");
        reserved = (bytes + gran - 1) / gran * gran;
        CUresult r = g_vmm.AddressReserve(&base, reserved, 0, 0, 0);
        if (r != CUDA_SUCCESS) { base = 0; return fail(-6, "cuMemAddressReserve(%zu) failed with CUresult %d", reserved, (int)r); }
        return 0;
    }
    // Make at least `bytes` from the base usable.
    int grow(size_t bytes) {
        if (bytes <= mapped) return 0;
        if (bytes > reserved) return fail(-6, "store array would exceed its reserved address range (%zu > %zu)", bytes, reserved);
        size_t want = (bytes - mapped + gran - 1) / gran * gran;
        CUmemAllocationProp prop;
        memset(&prop, 0, sizeof(prop));
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = device;
        CUmemGenericAllocationHandle h;
        CUresult r = g_vmm.Create(&h, want, &prop, 0);
        if (r != CUDA_SUCCESS) return fail(-6, "cuMemCreate(%zu) failed with CUresult %d (out of HBM?)", want, (int)r);
        r = g_vmm.Map(base + mapped, want, 0, h, 0);
        if (r != CUDA_SUCCESS) { g_vmm.Release(h); return fail(-6, "cuMemMap failed with CUresult %d", (int)r); }
        CUmemAccessDesc acc;
        memset(&acc, 0, sizeof(acc));
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = device;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        r = g_vmm.SetAccess(base + mapped, want, &acc, 1);
        if (r != CUDA_SUCCESS) { g_vmm.Unmap(base + mapped, want); g_vmm.Release(h); return fail(-6, "cuMemSetAccess failed with CUresult %d", (int)r); }
        chunks.emplace_back(h, want);
        mapped += want;
        return 0;
    }
    void release() {
        if (!base) return;
        size_t off = 0;
        for (auto& c : chunks) { g_vmm.Unmap(base + off, c.second); g_vmm.Release(c.first); off += c.second; }
        chunks.clear();
        g_vmm.AddressFree(base, reserved);
        base = 0; reserved = mapped = 0;
    }
};

// ---------------------------------------------------------------------------
// the store
// ---------------------------------------------------------------------------
struct tfp_find_ctx;
struct tfp_multi;

struct tfp_store {
    tfp_multi* multi = nullptr;   // set: this handle is ONE logical store over several GPUs (multi_store.inl); the fields below are unused
    std::mutex err_mu;
    std::string last_err;         // text of the last failing call on this handle, from any thread (tfp_store_last_error)
    int device = 0;
    std::mutex mut_mu;         // one mutation (set / delete / get / save / load) at a time: owns the staging buffers, the host index and its mirrors
    RwMutex rw;      // finds: shared while enqueueing.  Mutations take it exclusively only to PUBLISH (row count, flags,
                               // capacity, tensor maps) and, for overwrites and compaction, to fence the finds; taken after mut_mu
    int dim = -1;              // store.go:29,44 keyLen
    bool normalized = true;    // store.go:27,43 keysAreNormalized (sticky)
    int64_t n = 0, cap = 0, cap_hint = 0;
    int64_t next_id = 0;
    int64_t hbm_mapped = 0;    // bytes of HBM mapped behind the three arrays (published with cap)
    VmmArray vm_keys, vm_ids, vm_knorm;   // backing of the three arrays when virtual memory management is available
    float* d_keys = nullptr;
    int64_t* d_ids = nullptr;
    double* d_knorm = nullptr;
    CUtensorMap tmap;        // exact scan: box [32 floats x 128 rows], 128-byte swizzle
    CUtensorMap tmap_mma;    // tensor-core path: box [16 floats x 128 rows], 64-byte swizzle
    bool tmap_ok = false;
    FlatIndex index;
    std::vector<uint64_t> slot_hash;   // host mirror: slot -> row hash
    std::vector<int64_t> slot_id;      // host mirror: slot -> id
    cudaStream_t s_mut = nullptr;      // side stream for ingest / compaction
    // Tensor-core passes of different find contexts run one after the other (each uses every SM); what overlaps with
    // the next batch's passes is a batch's finalize.  mma_ev is recorded behind every full pass.
    std::mutex mma_mu;
    cudaEvent_t mma_ev = nullptr;
    tfp_find_ctx* mma_last = nullptr;
    DevBuf st_rows, st_hash, st_knorm, st_isnorm, st_pair_row, st_pair_slot, st_pair_eq, st_src, st_dst, st_ids;
    std::vector<tfp_find_ctx*> ctxs;
    tfp_find_ctx* default_ctx = nullptr;
    std::mutex find_mu;                // the built-in context serves one synchronous find at a time
    int sm_count = 148;
    int max_smem = 0;
    double tr_us[6] = {0, 0, 0, 0, 0, 0};   // TFP_SET_TRACE: hash + D2H, index probe, verify, grow / fence, scatter + sync, publish
    int64_t tr_calls = 0;
    double trd_us[5] = {0, 0, 0, 0, 0};     // ... of StoresDelete: sort + look-up, wait for the lock, wait for the finds, host index, move + sync
    int64_t trd_calls = 0;
};

struct tfp_find_ctx {
    tfp_store* st = nullptr;
    cudaStream_t stream = nullptr;
    int nq = 0, dim = 0;
    std::vector<QInfo> qinfo;
    DevBuf d_in;                         // staged queries, then their QInfo records (one H2D copy)
    float* d_q = nullptr;
    QInfo* d_qinfo = nullptr;
    PinBuf h_in, h_out;
    DevBuf part, part_cnt, sync_words;   // sync_words: gtau[kMaxNQ] + ticket + tile counter + lbest[kMaxNQ x kMaxGrid]
    DevBuf all_cand, sorted_cand, cub_tmp;
    DevBuf dbg_times; int dbg_grid = 0;
    DevBuf mma_qsplit, mma_cand, mma_small;   // tensor-core path: split queries, candidate lists, counters/bounds/status
    PinBuf h_status;
    // sharded finds: communicator, gathered records, merged answer
    void* nccl_comm = nullptr;
    bool comm_owned = false;
    int comm_rank = 0, comm_world = 0;
    DevBuf gathered, merged;
    int merged_topk = 0, merged_count = 0;
    int merged_keys_dim = 0;             // > 0: the merged block carries the winners' key rows after its fixed part
    // sharded finds over peer memory (tfp_find_ctx_p2p_*): this rank's gather block
    // [2 parities][world][x_slot bytes] + [world] sequence words, and every rank's block mapped here
    DevBuf xbuf;
    size_t x_slot = 0;
    uint8_t* x_peer[kXchgMaxWorld] = {nullptr};
    int x_world = 0, x_me = 0;
    unsigned int x_seq = 0;
    bool x_ready = false;
    bool x_root_mode = false;            // one process over several GPUs: records are gathered on shard 0 only
    int x_m_count = 0;                   // the pending peer-exchange merge: queries, and whether key rows travel too
    bool x_m_with_keys = false;
    bool x_no_spin = false;              // the merge must not wait inside the kernel (shards sharing one GPU): the caller orders it with events
    cudaEvent_t x_done_ev = nullptr;     // ... recorded on this context's stream after its scan and pushes
    bool x_pushed = false;               // the last launch pushed its records from the scan kernel itself
    bool keys_gathered = false;          // the last launch's scan kernels wrote the winning key rows themselves
    int64_t mma_passes = 0, mma_fallbacks = 0, mma_escalated = 0;
    // Results of the last launch live in ONE block so that a download is one D2H copy:
    // [n | ids | sims | key rows] first (what the host reads), then [cand | slots | records].
    DevBuf out_block;
    int* out_n = nullptr; int64_t* out_ids = nullptr; float* out_sims = nullptr; float* out_keys = nullptr;
    uint64_t* out_cand = nullptr; uint32_t* out_slots = nullptr; uint8_t* out_rec = nullptr;
    size_t off_ids = 0, off_sims = 0, off_keys = 0;   // byte offsets inside out_block
    int last_count = 0, last_topk = 0, last_stride = 0, last_path = 0;
    int last_topk_host = 0;              // the topk the caller passed (host arrays are strided by it; last_topk may be clamped)
    int64_t launches = 0;
    struct MmaPending* pend = nullptr;   // a tensor-core launch whose status words have not been read yet (ctx_resolve_locked)
};

static int set_device(const tfp_store* s) {
    CU_OK(cudaSetDevice(s->device));
    return 0;
}

static bool tma_eligible(const tfp_store* s) { return s->dim >= kBoxCols && (s->dim % 4) == 0; }

static int encode_tmap(tfp_store* s) {
    s->tmap_ok = false;
    if (!tma_eligible(s) || s->cap < kRowsPerTile) return 0;
    if (get_encoder()) return -2;
    cuuint64_t gdim[2] = {(cuuint64_t)s->dim, (cuuint64_t)s->cap};
    cuuint64_t gstride[1] = {(cuuint64_t)s->dim * 4};
    cuuint32_t box[2] = {(cuuint32_t)kBoxCols, (cuuint32_t)kRowsPerTile};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(&s->tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, s->d_keys, gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(-2, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    cuuint32_t box2[2] = {(cuuint32_t)kMmaBoxCols, (cuuint32_t)kMmaRows};
    r = g_encode(&s->tmap_mma, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, s->d_keys, gdim, gstride, box2, estr,
                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(-2, "cuTensorMapEncodeTiled (mma) failed with CUresult %d", (int)r);
    s->tmap_ok = true;
    return 0;
}

// Grow the HBM arrays to hold `rows` rows (dim must be latched).  Caller holds mut_mu.
// publish: the store lock to take -- exclusively, if the caller does not hold it yet -- for the instant in which what
// finds read when they launch (capacity, tensor maps, the first time the base addresses) changes.  With virtual
// memory management the slow part, creating and mapping physical memory (tens of milliseconds for a multi-GB step: the
// driver zeroes it), happens BEFORE that and stops no find: running kernels keep their own copy of the tensor map and
// never touch the new pages.
static int ensure_capacity(tfp_store* s, int64_t rows, std::unique_lock<RwMutex>* publish = nullptr) {
    if (rows <= s->cap) return 0;
    if (rows > 0xFFFFFFF0ll) return fail(-3, "store shard would exceed 2^32 rows");
    const size_t row_bytes = (size_t)s->dim * 4;
    auto lock_for_publish = [&]() { if (publish && !publish->owns_lock()) publish->lock(); };
    if (vmm_api()) {
        // virtual memory: map more physical memory behind fixed addresses; nothing is copied
        const bool first = !s->vm_keys.active();
        if (first) {
            size_t free_b = 0, total_b = 0;
            CU_OK(cudaMemGetInfo(&free_b, &total_b));
            int64_t max_rows = (int64_t)std::min<size_t>((size_t)0xFFFFFFF0ull, total_b / row_bytes + 1);
            if (s->vm_keys.reserve(s->device, (size_t)max_rows * row_bytes + 4096) ||
                s->vm_ids.reserve(s->device, (size_t)max_rows * 8) || s->vm_knorm.reserve(s->device, (size_t)max_rows * 8))
                return -6;
        }
        // a quarter more than there is, but no more than 2 GiB of keys at a time: the call that grows pays for the step
        const int64_t step = std::min<int64_t>(s->cap / 4, (int64_t)(((size_t)2 << 30) / row_bytes));
        int64_t ncap = std::max<int64_t>({rows, s->cap + step, s->cap_hint, (int64_t)1024});
        ncap = std::min<int64_t>(ncap, 0xFFFFFFF0ll);
        if (s->vm_keys.grow((size_t)ncap * row_bytes + 256) || s->vm_ids.grow((size_t)ncap * 8) || s->vm_knorm.grow((size_t)ncap * 8)) {
            // the generous step did not fit: take exactly what is needed
            ncap = rows;
            if (s->vm_keys.grow((size_t)ncap * row_bytes + 256) || s->vm_ids.grow((size_t)ncap * 8) || s->vm_knorm.grow((size_t)ncap * 8))
                return -6;
        }
        // whatever the mapping granularity added is usable too
        int64_t by_map = (int64_t)std::min({(s->vm_keys.mapped - 256) / row_bytes, s->vm_ids.mapped / 8, s->vm_knorm.mapped / 8});
        lock_for_publish();
        if (first) {
            s->d_keys = (float*)s->vm_keys.base;
            s->d_ids = (int64_t*)s->vm_ids.base;
            s->d_knorm = (double*)s->vm_knorm.base;
        }
        s->cap = std::min<int64_t>(std::max(ncap, by_map), 0xFFFFFFF0ll);
        s->hbm_mapped = (int64_t)(s->vm_keys.mapped + s->vm_ids.mapped + s->vm_knorm.mapped);
        return encode_tmap(s);
    }
    lock_for_publish();
    // no virtual memory management: reallocate and copy
    int64_t ncap = std::max<int64_t>({rows, s->cap * 2, s->cap_hint, (int64_t)1024});
    float* nk = nullptr; int64_t* ni = nullptr; double* nn = nullptr;
    auto grow = [&]() -> int {
        CU_OK(cudaMalloc(&nk, (size_t)ncap * s->dim * 4 + 256));
        CU_OK(cudaMalloc(&ni, (size_t)ncap * 8));
        CU_OK(cudaMalloc(&nn, (size_t)ncap * 8));
        if (s->n > 0) {
            CU_OK(cudaMemcpyAsync(nk, s->d_keys, (size_t)s->n * s->dim * 4, cudaMemcpyDeviceToDevice, s->s_mut));
            CU_OK(cudaMemcpyAsync(ni, s->d_ids, (size_t)s->n * 8, cudaMemcpyDeviceToDevice, s->s_mut));
            CU_OK(cudaMemcpyAsync(nn, s->d_knorm, (size_t)s->n * 8, cudaMemcpyDeviceToDevice, s->s_mut));
            CU_OK(cudaStreamSynchronize(s->s_mut));
        }
        return 0;
    };
    if (grow()) {   // the store keeps its old arrays; nothing half-grown is left behind
        if (nk) cudaFree(nk);
        if (ni) cudaFree(ni);
        if (nn) cudaFree(nn);
        return -1;
    }
    if (s->d_keys) cudaFree(s->d_keys);
    if (s->d_ids) cudaFree(s->d_ids);
    if (s->d_knorm) cudaFree(s->d_knorm);
    s->d_keys = nk; s->d_ids = ni; s->d_knorm = nn; s->cap = ncap;
    return encode_tmap(s);
}

// ---------------------------------------------------------------------------
// kernel attribute setup (dynamic shared memory opt-in), once per process+device
// ---------------------------------------------------------------------------
template <int NQ, bool F64, bool SPILL> static cudaError_t optin_tma(int bytes) {
    return cudaFuncSetAttribute(scan_tma_kernel<NQ, F64, SPILL>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
template <int NQ, bool F64, bool SPILL> static cudaError_t optin_gen(int bytes) {
    return cudaFuncSetAttribute(scan_generic_kernel<NQ, F64, SPILL>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
static int optin_all(int bytes) {
    CU_OK((optin_tma<1, false, false>(bytes))); CU_OK((optin_tma<1, true, false>(bytes)));
    CU_OK((optin_tma<2, false, false>(bytes))); CU_OK((optin_tma<2, true, false>(bytes)));
    CU_OK((optin_tma<4, false, false>(bytes))); CU_OK((optin_tma<4, true, false>(bytes)));
    CU_OK((optin_tma<8, false, false>(bytes))); CU_OK((optin_tma<8, true, false>(bytes)));
    CU_OK((optin_tma<1, false, true>(bytes)));  CU_OK((optin_tma<1, true, true>(bytes)));
    CU_OK((optin_gen<1, false, false>(bytes))); CU_OK((optin_gen<1, true, false>(bytes)));
    CU_OK((optin_gen<2, false, false>(bytes))); CU_OK((optin_gen<2, true, false>(bytes)));
    CU_OK((optin_gen<4, false, false>(bytes))); CU_OK((optin_gen<4, true, false>(bytes)));
    CU_OK((optin_gen<8, false, false>(bytes))); CU_OK((optin_gen<8, true, false>(bytes)));
    CU_OK((optin_gen<1, false, true>(bytes)));  CU_OK((optin_gen<1, true, true>(bytes)));
    CU_OK((cudaFuncSetAttribute(mma_scan_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)));
    CU_OK((cudaFuncSetAttribute(mma_scan_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)));
    CU_OK((cudaFuncSetAttribute(mma_scan_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)));
    CU_OK((cudaFuncSetAttribute(mma_finalize_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFinDynSmem)));
    CU_OK((cudaFuncSetAttribute(mma_finalize_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFinDynSmem)));
    CU_OK((cudaFuncSetAttribute(mma_finalize_kernel<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFinDynSmem)));
    CU_OK((cudaFuncSetAttribute(mma_finalize_kernel<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFinDynSmem)));
    return 0;
}

// ---------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------
static int ctx_create(tfp_store* s, tfp_find_ctx** out);
static void ctx_destroy(tfp_find_ctx* c);
static int ctx_resolve_locked(tfp_find_ctx* c);
static void ctx_drop_pending(tfp_find_ctx* c);

static int store_open(const tfp_cfg* cfg, tfp_store** out) {
    if (!out) return fail(2, "tfp_store_open: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(-1, "no usable CUDA device (%s); libtfpstore has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    tfp_store* s = new tfp_store();
    s->device = cfg ? cfg->device : 0;
    s->cap_hint = cfg ? cfg->capacity_rows : 0;
    s->next_id = cfg ? cfg->id_base : 0;
    if (s->device < 0 || s->device >= ndev) { delete s; return fail(2, "device %d out of range (have %d)", cfg->device, ndev); }
    if (set_device(s)) { delete s; return -1; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, s->device) != cudaSuccess) { delete s; return fail(-1, "cudaGetDeviceProperties failed"); }
    if (prop.major != 10) {
        int maj = prop.major, min = prop.minor;
        delete s;
        return fail(-1, "device %d is sm_%d%d; libtfpstore is built for sm_100a (B200) only", cfg ? cfg->device : 0, maj, min);
    }
    s->sm_count = prop.multiProcessorCount;
    s->max_smem = (int)prop.sharedMemPerBlockOptin;
    if (optin_all(s->max_smem)) { delete s; return -1; }
    {   // the mutation stream outranks the find streams: its small kernels slot in as scan CTAs retire instead of
        // queueing behind every find in flight
        int prio_lo = 0, prio_hi = 0;
        cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
        if (cudaStreamCreateWithPriority(&s->s_mut, cudaStreamNonBlocking, prio_hi) != cudaSuccess) { delete s; return fail(-1, "cudaStreamCreate failed"); }
    }
    if (ctx_create(s, &s->default_ctx)) { cudaStreamDestroy(s->s_mut); delete s; return -1; }
    *out = s;
    return 0;
}

static void store_close(tfp_store* s) {
    if (!s) return;
    if (tuning().set_trace && s->tr_calls > 0)
        fprintf(stderr, "tfp set trace (device %d, %lld calls, us per call): hash + D2H %.0f | index probe %.0f | verify %.0f | grow / fence %.0f | "
                        "scatter + sync %.0f | publish %.0f\n", s->device, (long long)s->tr_calls, s->tr_us[0] / s->tr_calls,
                s->tr_us[1] / s->tr_calls, s->tr_us[2] / s->tr_calls, s->tr_us[3] / s->tr_calls, s->tr_us[4] / s->tr_calls,
                s->tr_us[5] / s->tr_calls);
    if (tuning().set_trace && s->trd_calls > 0)
        fprintf(stderr, "tfp delete trace (device %d, %lld calls, us per call): sort + look-up %.0f | wait for the store lock %.0f | wait for the finds in flight %.0f | "
                        "host index %.0f | move + sync %.0f\n", s->device, (long long)s->trd_calls, s->trd_us[0] / s->trd_calls, s->trd_us[1] / s->trd_calls,
                s->trd_us[2] / s->trd_calls, s->trd_us[3] / s->trd_calls, s->trd_us[4] / s->trd_calls);
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    std::vector<tfp_find_ctx*> ctxs = s->ctxs;
    for (tfp_find_ctx* c : ctxs) ctx_destroy(c);
    if (s->vm_keys.active()) {
        s->vm_keys.release(); s->vm_ids.release(); s->vm_knorm.release();
    } else {
        if (s->d_keys) cudaFree(s->d_keys);
        if (s->d_ids) cudaFree(s->d_ids);
        if (s->d_knorm) cudaFree(s->d_knorm);
    }
    DevBuf* bufs[] = {&s->st_rows, &s->st_hash, &s->st_knorm, &s->st_isnorm, &s->st_pair_row,
                      &s->st_pair_slot, &s->st_pair_eq, &s->st_src, &s->st_dst, &s->st_ids};
    for (DevBuf* b : bufs) b->release();
    if (s->s_mut) cudaStreamDestroy(s->s_mut);
    if (s->mma_ev) cudaEventDestroy(s->mma_ev);
    delete s;
}

static int store_stats(tfp_store* s, tfp_stats* out) {
    if (!s || !out) return fail(2, "tfp_store_stats: NULL argument");
    std::shared_lock<RwMutex> lk(s->rw);
    out->rows = s->n;
    out->capacity_rows = s->cap;
    out->hbm_bytes = s->vm_keys.active() ? s->hbm_mapped : s->cap * ((int64_t)(s->dim > 0 ? s->dim : 0) * 4 + 16);
    out->dim = s->dim;
    out->keys_are_normalized = s->normalized ? 1 : 0;
    out->device = s->device;
    out->reserved = 0;
    return 0;
}

static int store_mark_not_normalized(tfp_store* s) {
    if (!s) return fail(2, "tfp_store_mark_not_normalized: store is NULL");
    std::unique_lock<RwMutex> lk(s->rw);
    s->normalized = false;
    return 0;
}

// ---------------------------------------------------------------------------
// mutation plumbing
// ---------------------------------------------------------------------------
// Caller holds the exclusive lock.  A context whose tensor-core batch still waits for its status words gets the
// rest of that find enqueued first: it must read the rows the first pass read, before they move.
static int quiesce_finds(tfp_store* s) {
    for (tfp_find_ctx* c : s->ctxs) {
        if (ctx_resolve_locked(c)) return -1;
        CU_OK(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

// Go's cmp.Compare on float32 (NaN lowest, NaN == NaN, -0 == +0), used to put a
// request's keys in the reference's sorted order (store.go:86-99).
static inline int go_cmp(float x, float y) {
    bool xn = x != x, yn = y != y;
    if (xn) return yn ? 0 : -1;
    if (yn) return 1;
    return x < y ? -1 : (x > y ? 1 : 0);
}
static int cmp_rows(const float* a, const float* b, int d) {
    for (int i = 0; i < d; i++) { int c = go_cmp(a[i], b[i]); if (c) return c; }
    return 0;
}

struct Lookup {
    std::vector<uint64_t> hash;     // per staged row
    std::vector<uint8_t> is_norm;
    std::vector<int64_t> match;     // stored slot confirmed equal and not yet consumed, or -1
};

static void sorted_order(const float* keys, int64_t n, int dim, std::vector<int64_t>& order);

// Hash the staged rows on the device, probe the host index, confirm candidates on
// the device.  Rows claim stored slots in a fixed sequence (each stored slot is
// claimed at most once per call, like the reference's walks over a sorted batch:
// store.go:154-189, :225-238, :285-301).  That sequence only matters when the request
// holds the same key more than once, so the reference's sorted order is worked out
// -- from `host_keys`, the request's rows in host memory -- only when two rows share a
// hash; otherwise rows claim in request order.  (Sorting every request cost more than
// everything else a StoresDelete does: 35 ms for 6.5k keys of 1536 floats, each
// comparison two cache misses.)
static inline double now_us() {
    return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int lookup_rows(tfp_store* s, const float* d_rows, int64_t n, const float* host_keys, Lookup& L) {
    const int dim = s->dim;
    const double t0 = now_us();
    if (s->st_hash.reserve((size_t)n * 8) || s->st_knorm.reserve((size_t)n * 8) || s->st_isnorm.reserve((size_t)n)) return -1;
    int threads = kStatRows;
    int64_t blocks = (n + threads - 1) / threads;
    row_stats_kernel<<<(unsigned)blocks, threads, 0, s->s_mut>>>(d_rows, n, dim, s->st_hash.as<uint64_t>(),
                                                               s->st_knorm.as<double>(), s->st_isnorm.as<uint8_t>());
    CU_OK(cudaGetLastError());
    L.hash.resize((size_t)n);
    L.is_norm.resize((size_t)n);
    CU_OK(cudaMemcpyAsync(L.hash.data(), s->st_hash.p, (size_t)n * 8, cudaMemcpyDeviceToHost, s->s_mut));
    CU_OK(cudaMemcpyAsync(L.is_norm.data(), s->st_isnorm.p, (size_t)n, cudaMemcpyDeviceToHost, s->s_mut));
    CU_OK(cudaStreamSynchronize(s->s_mut));
    const double t1 = now_us();
    s->tr_us[0] += t1 - t0;

    // candidate pairs (row, slot)
    std::vector<uint32_t> pr, ps;
    std::vector<size_t> first((size_t)n + 1, 0);
    for (int64_t i = 0; i < n; i++) {
        first[(size_t)i] = pr.size();
        s->index.for_each(L.hash[(size_t)i], [&](uint32_t slot) { pr.push_back((uint32_t)i); ps.push_back(slot); });
    }
    first[(size_t)n] = pr.size();
    const double t2 = now_us();
    s->tr_us[1] += t2 - t1;
    std::vector<uint8_t> eq(pr.size(), 0);
    if (!pr.empty()) {
        size_t m = pr.size();
        if (s->st_pair_row.reserve(m * 4) || s->st_pair_slot.reserve(m * 4) || s->st_pair_eq.reserve(m)) return -1;
        CU_OK(cudaMemcpyAsync(s->st_pair_row.p, pr.data(), m * 4, cudaMemcpyHostToDevice, s->s_mut));
        CU_OK(cudaMemcpyAsync(s->st_pair_slot.p, ps.data(), m * 4, cudaMemcpyHostToDevice, s->s_mut));
        int64_t vthreads = (int64_t)m * 32;
        verify_pairs_kernel<<<(unsigned)((vthreads + 255) / 256), 256, 0, s->s_mut>>>(
            s->d_keys, dim, d_rows, s->st_pair_row.as<uint32_t>(), s->st_pair_slot.as<uint32_t>(), (int64_t)m,
            s->st_pair_eq.as<uint8_t>());
        CU_OK(cudaGetLastError());
        CU_OK(cudaMemcpyAsync(eq.data(), s->st_pair_eq.p, m, cudaMemcpyDeviceToHost, s->s_mut));
        CU_OK(cudaStreamSynchronize(s->s_mut));
    }
    s->tr_us[2] += now_us() - t2;
    L.match.assign((size_t)n, -1);
    std::vector<int64_t> order;
    if (host_keys && n > 1) {
        std::vector<uint64_t> hs(L.hash);
        std::sort(hs.begin(), hs.end());
        if (std::adjacent_find(hs.begin(), hs.end()) != hs.end()) sorted_order(host_keys, n, dim, order);
    }
    std::unordered_set<uint32_t> claimed;
    for (int64_t oi = 0; oi < n; oi++) {
        int64_t i = order.empty() ? oi : order[(size_t)oi];
        for (size_t j = first[(size_t)i]; j < first[(size_t)i + 1]; j++) {
            if (eq[j] && !claimed.count(ps[j])) {
                claimed.insert(ps[j]);
                L.match[(size_t)i] = ps[j];
                break;
            }
        }
    }
    return 0;
}

// Stage host rows in HBM on the mutation stream.
static int stage_rows(tfp_store* s, const float* keys, int64_t n, int dim, const float** d_rows) {
    size_t bytes = (size_t)n * dim * 4;
    if (s->st_rows.reserve(bytes)) return -1;
    CU_OK(cudaMemcpyAsync(s->st_rows.p, keys, bytes, cudaMemcpyHostToDevice, s->s_mut));
    *d_rows = s->st_rows.as<float>();
    return 0;
}

// Caller holds mut_mu (not rw).  Pure appends never stop the finds: new rows land beyond the published row count,
// which a scan launched earlier never reads (a tile that straddles the old end loads them and masks them), HBM
// grows by mapping more memory behind the same addresses, and the new row count is published -- under the
// exclusive lock, for the instant that takes -- only after the rows are in place.  An overwrite of a stored row
// (store.go:183-187) or a grow that has to move the arrays (no virtual memory management) fences the finds first.
static int set_impl(tfp_store* s, const float* d_rows, int64_t n, int64_t* out_ids, uint8_t* out_was_new) {
    const int dim = s->dim;
    Lookup L;
    if (lookup_rows(s, d_rows, n, nullptr, L)) return -1;
    bool all_norm = true;                                  // sticky flag (store.go:125-135)
    for (int64_t i = 0; i < n; i++) if (!L.is_norm[(size_t)i]) { all_norm = false; break; }

    std::vector<uint32_t> src((size_t)n), dst((size_t)n);
    std::vector<int64_t> ids((size_t)n);
    int64_t n_new = 0;
    for (int64_t i = 0; i < n; i++) {
        src[(size_t)i] = (uint32_t)i;
        if (L.match[(size_t)i] >= 0) {   // equal stored key: the new pair replaces it (store.go:183-187)
            dst[(size_t)i] = (uint32_t)L.match[(size_t)i];
            ids[(size_t)i] = s->slot_id[(size_t)L.match[(size_t)i]];
            if (out_was_new) out_was_new[i] = 0;
        } else {
            dst[(size_t)i] = (uint32_t)(s->n + n_new);
            ids[(size_t)i] = s->next_id + n_new;
            n_new++;
            if (out_was_new) out_was_new[i] = 1;
        }
        if (out_ids) out_ids[i] = ids[(size_t)i];
    }
    const double t3 = now_us();
    const bool overwrites = n_new != n;
    const bool must_move = s->n + n_new > s->cap && !(vmm_api() && (s->vm_keys.active() || s->cap == 0));
    std::unique_lock<RwMutex> fence(s->rw, std::defer_lock);
    if (overwrites || must_move) {
        fence.lock();
        if (quiesce_finds(s)) return -1;
    }
    if (s->n + n_new > s->cap) {
        // growing changes what finds read when they launch; ensure_capacity takes the lock only to publish that
        std::unique_lock<RwMutex> grow(s->rw, std::defer_lock);
        if (ensure_capacity(s, s->n + n_new, fence.owns_lock() ? nullptr : &grow)) return -1;
    }
    const double t4 = now_us();
    s->tr_us[3] += t4 - t3;
    if (s->st_src.reserve((size_t)n * 4) || s->st_dst.reserve((size_t)n * 4) || s->st_ids.reserve((size_t)n * 8)) return -1;
    CU_OK(cudaMemcpyAsync(s->st_src.p, src.data(), (size_t)n * 4, cudaMemcpyHostToDevice, s->s_mut));
    CU_OK(cudaMemcpyAsync(s->st_dst.p, dst.data(), (size_t)n * 4, cudaMemcpyHostToDevice, s->s_mut));
    CU_OK(cudaMemcpyAsync(s->st_ids.p, ids.data(), (size_t)n * 8, cudaMemcpyHostToDevice, s->s_mut));
    int threads = dim >= 256 ? 256 : (dim >= 64 ? 64 : 32);
    scatter_rows_kernel<<<(unsigned)n, threads, 0, s->s_mut>>>(d_rows, s->st_knorm.as<double>(), s->st_src.as<uint32_t>(),
                                                             s->st_dst.as<uint32_t>(), s->st_ids.as<int64_t>(), n, dim,
                                                             s->d_keys, s->d_ids, s->d_knorm);
    CU_OK(cudaGetLastError());
    // host mirrors and index (mutation-side state: mut_mu)
    s->slot_hash.resize((size_t)(s->n + n_new));
    s->slot_id.resize((size_t)(s->n + n_new));
    s->index.reserve((size_t)(s->n + n_new));
    for (int64_t i = 0; i < n; i++) {
        if (L.match[(size_t)i] < 0) {
            uint32_t slot = dst[(size_t)i];
            s->slot_hash[slot] = L.hash[(size_t)i];
            s->slot_id[slot] = ids[(size_t)i];
            s->index.insert(L.hash[(size_t)i], slot);
        }
    }
    CU_OK(cudaStreamSynchronize(s->s_mut));   // the rows are in HBM ...
    const double t5 = now_us();
    s->tr_us[4] += t5 - t4;
    {                                         // ... now finds may see them
        std::unique_lock<RwMutex> pub(s->rw, std::defer_lock);
        if (!fence.owns_lock()) pub.lock();
        s->n += n_new;
        s->next_id += n_new;
        if (!all_norm) s->normalized = false;
    }
    s->tr_us[5] += now_us() - t5;
    s->tr_calls++;
    return 0;
}

// store.go:115-116 / :209-210 / :275-276: the first Set, Delete or Get fixes the key length.
static int latch_dim(tfp_store* s, int dim, const char* fmt) {
    {   // the usual case -- the length is fixed and matches -- needs no exclusive lock
        std::shared_lock<RwMutex> lk(s->rw);
        if (s->dim == dim) return 0;
    }
    std::unique_lock<RwMutex> lk(s->rw);
    if (s->dim == -1) s->dim = dim;
    else if (dim != s->dim) return fail(1, fmt, dim, s->dim);
    return 0;
}

static int check_scatter_grid(int64_t n) {
    if (n > 0x7FFFFFFFll) return fail(2, "too many keys in one call (%lld)", (long long)n);
    return 0;
}

static int store_set(tfp_store* s, const float* keys, int64_t n, int dim, int64_t n_values, int64_t* out_ids,
                     uint8_t* out_was_new) {
    if (!s) return fail(2, "tfp_store_set: store is NULL");
    if (n == 0) return fail(1, "no keys to add");                                     // store.go:107-109
    if (n != n_values) return fail(1, "len(keys) = %lld, len(values) = %lld", (long long)n, (long long)n_values);  // :111-113
    if (!keys || dim <= 0) return fail(2, "tfp_store_set: keys is NULL or dim <= 0");
    if (check_scatter_grid(n)) return 2;
    std::lock_guard<std::mutex> mut(s->mut_mu);
    if (latch_dim(s, dim, "Try to add key with length %d when existing length is %d")) return 1;   // :115-120
    if (set_device(s)) return -1;
    const float* d_rows = nullptr;
    if (stage_rows(s, keys, n, dim, &d_rows)) return -1;
    return set_impl(s, d_rows, n, out_ids, out_was_new);
}

static int store_set_device(tfp_store* s, const float* d_keys, int64_t n, int dim, int64_t* out_ids,
                            uint8_t* out_was_new) {
    if (!s) return fail(2, "tfp_store_set_device: store is NULL");
    if (n == 0) return fail(1, "no keys to add");
    if (!d_keys || dim <= 0) return fail(2, "tfp_store_set_device: d_keys is NULL or dim <= 0");
    if (check_scatter_grid(n)) return 2;
    std::lock_guard<std::mutex> mut(s->mut_mu);
    if (latch_dim(s, dim, "Try to add key with length %d when existing length is %d")) return 1;
    if (set_device(s)) return -1;
    CU_OK(cudaDeviceSynchronize());   // the caller's producer stream is unknown: order after all of it
    return set_impl(s, d_keys, n, out_ids, out_was_new);
}

// Request keys in the reference's sorted order (store.go:86-99); ties keep request order.
static void sorted_order(const float* keys, int64_t n, int dim, std::vector<int64_t>& order) {
    order.resize((size_t)n);
    for (int64_t i = 0; i < n; i++) order[(size_t)i] = i;
    if (n < 2) return;
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) {
        return cmp_rows(keys + (size_t)a * dim, keys + (size_t)b * dim, dim) < 0;
    });
}

static int store_get(tfp_store* s, const float* keys, int64_t n, int dim, int64_t* out_ids, int64_t* out_order,
                     int64_t* out_n_found) {
    if (!s) return fail(2, "tfp_store_get: store is NULL");
    if (out_n_found) *out_n_found = 0;
    if (n == 0) return fail(1, "no keys to get");   // reference panics here (store.go:276); documented deviation
    if (!keys || dim <= 0 || !out_ids) return fail(2, "tfp_store_get: NULL argument or dim <= 0");
    std::lock_guard<std::mutex> mut(s->mut_mu);      // uses the mutation staging buffers and the host index; finds go on
    if (latch_dim(s, dim, "Try to get a key with length %d when existing length is %d")) return 1;   // :275-280
    if (set_device(s)) return -1;
    for (int64_t i = 0; i < n; i++) out_ids[i] = -1;
    int64_t found = 0;
    if (s->n > 0) {
        const float* d_rows = nullptr;
        if (stage_rows(s, keys, n, dim, &d_rows)) return -1;
        Lookup L;
        if (lookup_rows(s, d_rows, n, keys, L)) return -1;
        std::vector<int64_t> hit;
        for (int64_t i = 0; i < n; i++)
            if (L.match[(size_t)i] >= 0) { out_ids[i] = s->slot_id[(size_t)L.match[(size_t)i]]; hit.push_back(i); }
        found = (int64_t)hit.size();
        if (out_order) {   // the found keys in the reference's sorted-key order (store.go:269,285-301); equal keys keep request order
            std::stable_sort(hit.begin(), hit.end(), [&](int64_t a, int64_t b) {
                return cmp_rows(keys + (size_t)a * dim, keys + (size_t)b * dim, dim) < 0;
            });
            for (int64_t j = 0; j < found; j++) out_order[j] = hit[(size_t)j];
        }
    }
    if (out_n_found) *out_n_found = found;
    return 0;
}

static int store_delete(tfp_store* s, const float* keys, int64_t n, int dim, int64_t* out_ids) {
    if (!s) return fail(2, "tfp_store_delete: store is NULL");
    if (n == 0) return fail(1, "no keys to delete");                                 // store.go:201-203
    if (!keys || dim <= 0) return fail(2, "tfp_store_delete: keys is NULL or dim <= 0");
    std::lock_guard<std::mutex> mut(s->mut_mu);
    if (latch_dim(s, dim, "Trying to delete key with length %d when existing length is %d")) return 1;   // :209-214
    if (out_ids) for (int64_t i = 0; i < n; i++) out_ids[i] = -1;
    if (s->n == 0) return 0;
    if (set_device(s)) return -1;
    const double td0 = now_us();
    const double saved_set[3] = {s->tr_us[0], s->tr_us[1], s->tr_us[2]};   // (the look-up below also feeds the set trace: undo that)
    const float* d_rows = nullptr;
    if (stage_rows(s, keys, n, dim, &d_rows)) return -1;
    Lookup L;
    if (lookup_rows(s, d_rows, n, keys, L)) return -1;
    std::vector<uint32_t> holes;
    for (int64_t i = 0; i < n; i++) {
        if (L.match[(size_t)i] >= 0) {
            uint32_t slot = (uint32_t)L.match[(size_t)i];
            holes.push_back(slot);
            if (out_ids) out_ids[i] = s->slot_id[slot];
        }
    }
    s->tr_us[0] = saved_set[0]; s->tr_us[1] = saved_set[1]; s->tr_us[2] = saved_set[2];
    if (holes.empty()) return 0;
    // compaction moves rows that running scans read, and shrinks the row count: fence the finds (the look-up above did not)
    const double td1 = now_us();
    std::unique_lock<RwMutex> lk(s->rw);
    const double td2 = now_us();
    if (quiesce_finds(s)) return -1;
    const double td3 = now_us();
    const int64_t new_n = s->n - (int64_t)holes.size();
    std::vector<uint8_t> dead((size_t)(s->n - new_n), 0);   // marks deleted slots in the tail [new_n, n)
    for (uint32_t h : holes) {
        s->index.erase(s->slot_hash[h], h);
        if ((int64_t)h >= new_n) dead[(size_t)(h - new_n)] = 1;
    }
    // survivors in the tail fill the holes below new_n
    std::vector<uint32_t> src, dst;
    int64_t t = new_n;
    for (uint32_t h : holes) {
        if ((int64_t)h >= new_n) continue;
        while (dead[(size_t)(t - new_n)]) t++;
        src.push_back((uint32_t)t);
        dst.push_back(h);
        s->index.relocate(s->slot_hash[(size_t)t], (uint32_t)t, h);
        s->slot_hash[h] = s->slot_hash[(size_t)t];
        s->slot_id[h] = s->slot_id[(size_t)t];
        t++;
    }
    const double td4 = now_us();
    if (!src.empty()) {
        size_t m = src.size();
        if (s->st_src.reserve(m * 4) || s->st_dst.reserve(m * 4)) return -1;
        CU_OK(cudaMemcpyAsync(s->st_src.p, src.data(), m * 4, cudaMemcpyHostToDevice, s->s_mut));
        CU_OK(cudaMemcpyAsync(s->st_dst.p, dst.data(), m * 4, cudaMemcpyHostToDevice, s->s_mut));
        int threads = s->dim >= 256 ? 256 : (s->dim >= 64 ? 64 : 32);
        move_rows_kernel<<<(unsigned)m, threads, 0, s->s_mut>>>(s->d_keys, s->d_ids, s->d_knorm, s->st_src.as<uint32_t>(),
                                                              s->st_dst.as<uint32_t>(), (int64_t)m, s->dim);
        CU_OK(cudaGetLastError());
        CU_OK(cudaStreamSynchronize(s->s_mut));
    }
    s->slot_hash.resize((size_t)new_n);
    s->slot_id.resize((size_t)new_n);
    s->n = new_n;
    s->trd_us[0] += td1 - td0; s->trd_us[1] += td2 - td1; s->trd_us[2] += td3 - td2; s->trd_us[3] += td4 - td3; s->trd_us[4] += now_us() - td4;
    s->trd_calls++;
    return 0;
}

// ---------------------------------------------------------------------------
// snapshot / restore
// ---------------------------------------------------------------------------
struct SnapHeader {
    char magic[8];        // "TFPSTOR2" ("TFPSTOR1": the same without the tag)
    int32_t dim;
    int32_t normalized;
    int64_t rows;
    int64_t next_id;
    uint64_t tag;         // the host's generation stamp: names the values file that belongs to this snapshot
};
constexpr size_t kSnapHeaderV1Bytes = 32;

// Write a file so that a crash at any moment leaves either the previous file or the complete new one:
// `path.tmp` is written and fsync'ed, then renamed over `path` (and the directory entry synced).
template <class Body>
static int write_file_atomically(const char* path, Body&& body) {
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(2, "tfp_store_save: cannot open %s", tmp.c_str());
    bool ok = body(f);
    ok = (fflush(f) == 0) && ok;
    ok = (fsync(fileno(f)) == 0) && ok;
    ok = (fclose(f) == 0) && ok;
    if (!ok) { remove(tmp.c_str()); return fail(2, "tfp_store_save: short write to %s", tmp.c_str()); }
    if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return fail(2, "tfp_store_save: cannot rename %s to %s", tmp.c_str(), path); }
    std::string dir(path);
    const size_t slash = dir.find_last_of('/');
    dir = slash == std::string::npos ? "." : dir.substr(0, slash ? slash : 1);
    const int dfd = open(dir.c_str(), O_RDONLY | O_DIRECTORY);
    if (dfd >= 0) { fsync(dfd); close(dfd); }
    return 0;
}

static int store_save(tfp_store* s, const char* path, uint64_t tag) {
    if (!s || !path) return fail(2, "tfp_store_save: NULL argument");
    std::lock_guard<std::mutex> mut(s->mut_mu);
    std::shared_lock<RwMutex> lk(s->rw);   // finds only read: they may go on while the rows are copied out
    if (set_device(s)) return -1;
    SnapHeader h;
    memcpy(h.magic, "TFPSTOR2", 8);
    h.dim = s->dim; h.normalized = s->normalized ? 1 : 0; h.rows = s->n; h.next_id = s->next_id; h.tag = tag;
    const size_t chunk = (size_t)64 << 20;
    PinBuf pin;
    if (s->n > 0 && pin.reserve(chunk)) return -1;
    int cuda_rc = 0;
    int rc = write_file_atomically(path, [&](FILE* f) -> bool {
        bool ok = fwrite(&h, sizeof(h), 1, f) == 1;
        auto dump = [&](const void* dptr, size_t bytes) {
            for (size_t off = 0; ok && !cuda_rc && off < bytes; off += chunk) {
                size_t m = std::min(chunk, bytes - off);
                if (cudaMemcpy(pin.p, (const uint8_t*)dptr + off, m, cudaMemcpyDeviceToHost) != cudaSuccess) { cuda_rc = -1; return; }
                ok = fwrite(pin.p, 1, m, f) == m;
            }
        };
        if (s->n > 0) { dump(s->d_keys, (size_t)s->n * s->dim * 4); dump(s->d_ids, (size_t)s->n * 8); }
        return ok && !cuda_rc;
    });
    pin.release();
    if (cuda_rc) return fail(-1, "tfp_store_save: device-to-host copy failed");
    return rc;
}

static int store_load(tfp_store* s, const char* path) {
    if (!s || !path) return fail(2, "tfp_store_load: NULL argument");
    std::lock_guard<std::mutex> mut(s->mut_mu);
    std::unique_lock<RwMutex> lk(s->rw);
    if (s->n != 0 || s->dim != -1) return fail(2, "tfp_store_load: the store is not empty");
    if (set_device(s)) return -1;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(2, "tfp_store_load: cannot open %s", path);
    SnapHeader h;
    memset(&h, 0, sizeof(h));
    bool hdr = fread(&h, kSnapHeaderV1Bytes, 1, f) == 1;
    if (hdr && memcmp(h.magic, "TFPSTOR2", 8) == 0) hdr = fread(&h.tag, sizeof(h.tag), 1, f) == 1;
    else if (hdr && memcmp(h.magic, "TFPSTOR1", 8) != 0) hdr = false;
    if (!hdr || h.rows < 0 || h.rows > 0xFFFFFFF0ll || (h.rows > 0 && h.dim <= 0)) {
        fclose(f);
        return fail(2, "tfp_store_load: %s is not a store snapshot", path);
    }
    s->dim = h.dim;
    s->next_id = h.next_id;
    s->normalized = h.normalized != 0;
    if (h.rows == 0) { fclose(f); return 0; }
    // any failure below leaves the store as empty as it was (its arrays may have grown; that is harmless)
    auto undo = [&]() { s->dim = -1; s->normalized = true; s->next_id = 0; s->n = 0; s->index = FlatIndex(); s->slot_hash.clear(); s->slot_id.clear(); };
    if (ensure_capacity(s, h.rows)) { fclose(f); undo(); return -1; }
    const size_t chunk = (size_t)64 << 20;
    PinBuf pin;
    if (pin.reserve(chunk)) { fclose(f); undo(); return -1; }
    bool ok = true;
    auto fill = [&](void* dptr, size_t bytes) -> int {
        for (size_t off = 0; ok && off < bytes; off += chunk) {
            size_t m = std::min(chunk, bytes - off);
            ok = fread(pin.p, 1, m, f) == m;
            if (ok) CU_OK(cudaMemcpy((uint8_t*)dptr + off, pin.p, m, cudaMemcpyHostToDevice));
        }
        return 0;
    };
    int rc = fill(s->d_keys, (size_t)h.rows * h.dim * 4);
    if (!rc) rc = fill(s->d_ids, (size_t)h.rows * 8);
    fclose(f);
    pin.release();
    if (rc) { undo(); return rc; }
    if (!ok) { undo(); return fail(2, "tfp_store_load: %s is truncated", path); }
    // recompute row hashes and norms on the device, rebuild the host index and mirrors
    const int64_t n = h.rows;
    auto rebuild = [&]() -> int {
        if (s->st_hash.reserve((size_t)n * 8) || s->st_isnorm.reserve((size_t)n)) return -1;
        row_stats_kernel<<<(unsigned)((n + kStatRows - 1) / kStatRows), kStatRows, 0, s->s_mut>>>(s->d_keys, n, s->dim, s->st_hash.as<uint64_t>(),
                                                                            s->d_knorm, s->st_isnorm.as<uint8_t>());
        CU_OK(cudaGetLastError());
        s->slot_hash.resize((size_t)n);
        s->slot_id.resize((size_t)n);
        CU_OK(cudaMemcpyAsync(s->slot_hash.data(), s->st_hash.p, (size_t)n * 8, cudaMemcpyDeviceToHost, s->s_mut));
        CU_OK(cudaMemcpyAsync(s->slot_id.data(), s->d_ids, (size_t)n * 8, cudaMemcpyDeviceToHost, s->s_mut));
        CU_OK(cudaStreamSynchronize(s->s_mut));
        return 0;
    };
    if (rebuild()) { undo(); return -1; }
    s->index.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++) s->index.insert(s->slot_hash[(size_t)i], (uint32_t)i);
    s->n = n;
    return 0;
}

// ---------------------------------------------------------------------------
// find contexts
// ---------------------------------------------------------------------------
static int ctx_create(tfp_store* s, tfp_find_ctx** out) {
    tfp_find_ctx* c = new tfp_find_ctx();
    c->st = s;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return fail(-1, "cudaStreamCreate failed"); }
    if (c->sync_words.reserve((kMaxNQ + 2 + kMaxNQ * kMaxGrid) * 8)) { cudaStreamDestroy(c->stream); delete c; return -1; }
    if (cudaMemset(c->sync_words.p, 0, (kMaxNQ + 2 + kMaxNQ * kMaxGrid) * 8) != cudaSuccess) { ctx_destroy(c); return fail(-1, "cudaMemset failed"); }
    s->ctxs.push_back(c);
    *out = c;
    return 0;
}
static void ctx_destroy(tfp_find_ctx* c) {
    if (!c) return;
    tfp_store* s = c->st;
    if (c->stream) cudaStreamSynchronize(c->stream);
    ctx_drop_pending(c);
    if (s) { std::lock_guard<std::mutex> g(s->mma_mu); if (s->mma_last == c) s->mma_last = nullptr; }
    if (c->nccl_comm && c->comm_owned && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl_comm);
    c->nccl_comm = nullptr;
    for (int r = 0; r < kXchgMaxWorld; r++)
        if (c->x_peer[r] && r != c->x_me && !c->x_root_mode) cudaIpcCloseMemHandle(c->x_peer[r]);
    c->xbuf.release();
    if (c->x_done_ev) cudaEventDestroy(c->x_done_ev);
    if (c->stream) cudaStreamDestroy(c->stream);
    DevBuf* bufs[] = {&c->d_in, &c->part, &c->part_cnt, &c->sync_words, &c->all_cand, &c->sorted_cand,
                      &c->cub_tmp, &c->out_block, &c->dbg_times, &c->mma_qsplit, &c->mma_cand, &c->mma_small, &c->gathered, &c->merged};
    for (DevBuf* b : bufs) b->release();
    c->h_in.release();
    c->h_out.release();
    c->h_status.release();
    if (s) s->ctxs.erase(std::remove(s->ctxs.begin(), s->ctxs.end(), c), s->ctxs.end());
    delete c;
}

extern "C" tfp_find_ctx* tfp_store_default_ctx(tfp_store* s) { return (s && !s->multi) ? s->default_ctx : nullptr; }

extern "C" int tfp_find_ctx_create(tfp_store* s, tfp_find_ctx** out) {
    if (!s || !out) return fail(2, "tfp_find_ctx_create: NULL argument");
    if (s->multi)
        return fail(2, "tfp_find_ctx_create: a multi-GPU store runs its own find lanes; call tfp_store_find / "
                       "tfp_store_find_batch from several threads instead");
    std::unique_lock<RwMutex> lk(s->rw);
    if (set_device(s)) return -1;
    return ctx_create(s, out);
}
extern "C" void tfp_find_ctx_destroy(tfp_find_ctx* c) {
    if (!c) return;
    tfp_store* s = c->st;
    std::unique_lock<RwMutex> lk(s->rw);
    cudaSetDevice(s->device);
    if (c == s->default_ctx) return;   // owned by the store
    ctx_destroy(c);
}

// store.go:313-324
static bool host_is_normalized(const float* k, int d) {
    double sum = 0.0;
    for (int i = 0; i < d; i++) { double v = (double)k[i]; sum += v * v; }
    double s = std::sqrt(sum);
    return s >= 0.99 && s <= 1.01;
}
// store.go:435-439
static double host_query_mag(const float* q, int d) {
    double m = 0.0;
    for (int i = 0; i < d; i++) { volatile float p = q[i] * q[i]; m += (double)p; }
    return std::sqrt(m);
}

// pre: the queries' QInfo records when the caller has computed them already (one logical store over several
// GPUs computes them once for all its shards), else nullptr.
static int ctx_upload_locked(tfp_find_ctx* c, const float* q, int nq, int dim, const QInfo* pre = nullptr) {
    tfp_store* s = c->st;
    size_t qbytes = (size_t)nq * dim * 4, ibytes = (size_t)nq * sizeof(QInfo);
    qbytes = (qbytes + 15) & ~(size_t)15;
    if (c->h_in.reserve(qbytes + ibytes) || c->d_in.reserve(qbytes + ibytes + 256)) return -1;
    c->d_q = c->d_in.as<float>();
    c->d_qinfo = (QInfo*)((uint8_t*)c->d_in.p + qbytes);
    c->qinfo.resize((size_t)nq);
    for (int i = 0; i < nq; i++) {
        const float* qi = q + (size_t)i * dim;
        QInfo& info = c->qinfo[(size_t)i];
        if (pre) { info = pre[i]; continue; }
        info.normalized = host_is_normalized(qi, dim) ? 1 : 0;   // the store flag is folded in at launch
        info.mag1 = host_query_mag(qi, dim);
        info.pad = 0;
    }
    memcpy(c->h_in.p, q, (size_t)nq * dim * 4);
    memcpy((uint8_t*)c->h_in.p + qbytes, c->qinfo.data(), ibytes);
    CU_OK(cudaMemcpyAsync(c->d_in.p, c->h_in.p, qbytes + ibytes, cudaMemcpyHostToDevice, c->stream));
    c->nq = nq;
    c->dim = dim;
    (void)s;
    return 0;
}

extern "C" int tfp_find_ctx_upload(tfp_find_ctx* c, const float* q, int nq, int dim) {
    if (!c || !q || nq <= 0 || dim <= 0) return fail(2, "tfp_find_ctx_upload: bad argument");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    if (int rc = ctx_resolve_locked(c)) return rc;   // a re-run would read the queries this call replaces
    CU_OK(cudaStreamSynchronize(c->stream));   // the pinned staging buffer is reused
    return ctx_upload_locked(c, q, nq, dim);
}

static int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

template <int NQ, bool F64, bool SPILL>
static int launch_scan(tfp_find_ctx* c, const ScanParams& p, bool use_tma, int grid, size_t smem) {
    tfp_store* s = c->st;
    if (use_tma) scan_tma_kernel<NQ, F64, SPILL><<<grid, kScanThreads, smem, c->stream>>>(s->tmap, p);
    else scan_generic_kernel<NQ, F64, SPILL><<<grid, kSelThreads, smem, c->stream>>>(p);
    CU_OK(cudaGetLastError());
    c->launches++;
    return 0;
}
template <bool F64, bool SPILL>
static int launch_scan_nq(tfp_find_ctx* c, int nq, const ScanParams& p, bool use_tma, int grid, size_t smem) {
    if constexpr (SPILL) { (void)nq; return launch_scan<1, F64, true>(c, p, use_tma, grid, smem); }
    else {
        switch (nq) {
            case 1: return launch_scan<1, F64, false>(c, p, use_tma, grid, smem);
            case 2: return launch_scan<2, F64, false>(c, p, use_tma, grid, smem);
            case 4: return launch_scan<4, F64, false>(c, p, use_tma, grid, smem);
            case 8: return launch_scan<8, F64, false>(c, p, use_tma, grid, smem);
        }
        return fail(-4, "internal: bad query group size %d", nq);
    }
}

// What one launch worked out about the store and the request: kept by value, so that the part of a tensor-core launch
// that runs after its status words arrive (ctx_resolve_locked) scans exactly the rows and slots the first pass did.
struct LaunchPlan {
    ScanParams p;
    bool fused = false;
    int dpad = 0, cap = 0, grid = 0, ksel = 0, count = 0, first = 0;
    uint32_t n_tiles = 0;
    int64_t n = 0;
};

// Launch the exact scan for staged queries [qf, qf+g) (same path, g in {1,2,4,8} or any g: it is
// reduced to what fits) writing output slots [of, ...).  Returns the group size actually launched.
// sel != nullptr: the g queries are sel[0..g) (any staged queries of this launch's range) and their answers go to
// their own slots, of + (sel[j] - qf); such launches do not push records to peers (the caller does, afterwards).
static int exact_launch(tfp_find_ctx* c, LaunchPlan& L, int qf, int of, int g, bool norm, int* launched, const int* sel = nullptr) {
    tfp_store* s = c->st;
    // pick the kernel flavour and the stage count that fits shared memory
    bool use_tma = s->tmap_ok;
    int n_stages = 0;
    size_t smem = 0;
    for (;;) {
        if (use_tma) {
            size_t fixed = scan_smem_bytes(0, g, L.dpad, L.cap) + 2 * kMaxStages * 8;
            int room = ((int)s->max_smem - (int)fixed) / kStageBytes;
            n_stages = std::min(kMaxStages, room);
            // A small shard (<= 32 tiles per CTA) scanned from a caller-made find context -- finds that are meant to
            // overlap -- gets a 6-stage ring: two CTAs then share an SM, so one find's prologue / publish / merge
            // runs under the other's streaming, and finds that run side by side read each tile from HBM once and
            // from L2 after that.  Measured on 125k x 768 (profiles/r02_probe_small_stages.log): 58 -> 51 us per find
            // with four finds in flight, 57 -> 40 with two; a lone scan pays 2-3 % (68.0 -> 69.3 us), so the
            // store's own synchronous context and large stores keep the deep ring.
            if (tuning().scan_stages >= 3) n_stages = std::min(n_stages, tuning().scan_stages);
            else if (tuning().scan_stages == 0 && c != s->default_ctx && L.n_tiles <= 32u * (uint32_t)L.grid)
                n_stages = std::min(n_stages, 6);
            if (n_stages >= 3) { smem = scan_smem_bytes(n_stages, g, L.dpad, L.cap); break; }
            if (g > 1) { g >>= 1; continue; }
            use_tma = false;
        }
        smem = scan_generic_smem_bytes(g, L.cap);
        if (smem <= (size_t)s->max_smem) break;
        if (g > 1) { g >>= 1; continue; }
        return fail(-4, "internal: selection buffers do not fit shared memory (k=%d)", L.ksel);
    }
    L.p.n_stages = n_stages;
    L.p.q_first = qf;
    L.p.out_first = of;
    L.p.use_qsel = sel ? 1 : 0;
    for (int j = 0; j < kMaxNQ; j++) L.p.qsel[j] = sel ? sel[j < g ? j : g - 1] : 0;
    L.p.xchg.raise = (of + g == L.count) ? 1 : 0;   // the launch that finishes the find raises the peers' words
    int rc;
    if (L.fused) rc = norm ? launch_scan_nq<false, false>(c, g, L.p, use_tma, L.grid, smem)
                         : launch_scan_nq<true, false>(c, g, L.p, use_tma, L.grid, smem);
    else       rc = norm ? launch_scan_nq<false, true>(c, 1, L.p, use_tma, L.grid, smem)
                         : launch_scan_nq<true, true>(c, 1, L.p, use_tma, L.grid, smem);
    if (rc) return rc;
    if (!L.fused) {
        // the spill kernels have no epilogue: re-arm the tile scheduler from the stream
        CU_OK(cudaMemsetAsync(L.p.tile_counter, 0, sizeof(unsigned int), c->stream));
        // big-k path: sort every row's candidate word (descending), keep the first k
        size_t tmp = 0;
        CU_OK(cub::DeviceRadixSort::SortKeysDescending(nullptr, tmp, c->all_cand.as<uint64_t>(),
                                                       c->sorted_cand.as<uint64_t>(), (int64_t)L.n, 0, 64, c->stream));
        if (c->cub_tmp.reserve(tmp)) return -1;
        CU_OK(cub::DeviceRadixSort::SortKeysDescending(c->cub_tmp.p, tmp, c->all_cand.as<uint64_t>(),
                                                       c->sorted_cand.as<uint64_t>(), (int64_t)L.n, 0, 64, c->stream));
        decode_sorted_kernel<<<std::min(1024, (L.ksel + 255) / 256), 256, 0, c->stream>>>(
            c->sorted_cand.as<uint64_t>(), (uint32_t)L.n, L.ksel, s->d_ids, of, L.p.out_cand, L.p.out_ids, L.p.out_sims,
            L.p.out_slots, L.p.out_n, L.p.out_rec, L.p.rec_bytes, L.p.rec_topk);
        CU_OK(cudaGetLastError());
        c->launches += 2;
    }
    if (launched) *launched = g;
    return 0;
}

// One tensor-core pass over the keys for g same-path staged queries named by index: h_qidx[j] is a staged query,
// its answer goes to output slot o_base + (h_qidx[j] - q_base).  Enqueues query split, bound-only pass, full pass
// and finalize with the `terms`-TF32 filter (1 or 3), then the copy of the per-query status words to h_status.
static int mma_pass_launch(tfp_find_ctx* c, const ScanParams& sp, const int* h_qidx, int g, int q_base, int o_base, bool norm,
                           int grid, int terms) {
    tfp_store* s = c->st;
    const int dim = s->dim;
    const int nqp = (g + 15) & ~15;
    const size_t small_bytes = (size_t)kMmaMaxNQ * 4 * 5 + (size_t)kMaxGrid * kMmaMaxNQ * 4;
    if (c->mma_qsplit.reserve((size_t)2 * kMmaMaxNQ * dim * 4 + 256) || c->mma_cand.reserve((size_t)kMmaMaxNQ * kMmaCandCap * 8) ||
        c->mma_small.reserve(small_bytes) || c->h_status.reserve(kMmaMaxNQ * 12)) return -1;
    unsigned int* cand_cnt = c->mma_small.as<unsigned int>();     // cand_cnt and gbound are adjacent: one memset per batch
    unsigned int* gbound = cand_cnt + kMmaMaxNQ;
    int* status = (int*)(gbound + kMmaMaxNQ);
    float* eps = (float*)(status + kMmaMaxNQ);
    int* qmap = (int*)(eps + kMmaMaxNQ);
    unsigned int* lbest2 = (unsigned int*)(qmap + kMmaMaxNQ);
    float* qs = c->mma_qsplit.as<float>();
    // pinned staging: [status | eps | qmap], one H2D copy for eps + qmap
    float* he = c->h_status.as<float>() + kMmaMaxNQ;
    int* hq = (int*)(he + kMmaMaxNQ);
    // Single-TF32 filter: |approx - exact| <= eps = unit |q| |k| with
    //   unit = 2^-9 + 2^-20   the operands lose < 2^-10 of their value each (TF32 keeps 10 mantissa bits), Cauchy-Schwarz
    //        + 2 D 2^-23      the tensor core's fp32 accumulation (adds that may truncate: D 2^-23) and the reference's own
    //                         sequential fp32 rounding (D 2^-24), rounded up
    // Normalised path: |q|, |k| <= 1.01.  Cosine path: the approximate score is dot / |k|, so the bound scales with
    // |q| only (+ 2^-22 for the fp32 reciprocal and multiply).  2 % slack on top.
    const float unit = (1.0f / 512.0f + 1.0f / 1048576.0f + 2.0f * (float)dim / 8388608.0f + 1.0f / 4194304.0f) * 1.02f;
    for (int j = 0; j < g; j++) {
        he[j] = terms == 1 ? (norm ? unit * 1.0201f : unit * (float)c->qinfo[(size_t)h_qidx[j]].mag1) : 0.0f;
        hq[j] = h_qidx[j];
    }
    for (int j = g; j < kMmaMaxNQ; j++) { he[j] = 0.0f; hq[j] = h_qidx[0]; }
    CU_OK(cudaMemcpyAsync(eps, he, (size_t)2 * kMmaMaxNQ * 4, cudaMemcpyHostToDevice, c->stream));
    // Two contexts' passes interleaved CTA by CTA can settle into a bad schedule on a large store: each grid holds part
    // of the SMs and both run at the pace of their late starters (1M x 768, 256 queries, two contexts in turn: 0.76 ms
    // per batch in most runs, 1.1 - 1.2 ms in one run out of three; profiles/r02_probe_defer.log).  So on 2 GiB of keys
    // or more this context's passes start behind the last full pass another context enqueued; what still overlaps is a
    // batch's finalize with the next batch's passes (0.78 - 0.90 ms there, every run).  Smaller stores, whose passes
    // do not fill the GPU for long, gain more from running freely (250k x 768, 64 queries: 0.167 ms against 0.197).
    const bool serial = tuning().mma_serial > 0 || (tuning().mma_serial < 0 && (size_t)sp.n_rows * dim * 4 >= ((size_t)2 << 30));
    if (serial) {
        std::lock_guard<std::mutex> g(s->mma_mu);
        if (!s->mma_ev) CU_OK(cudaEventCreateWithFlags(&s->mma_ev, cudaEventDisableTiming));
        if (s->mma_last && s->mma_last != c) CU_OK(cudaStreamWaitEvent(c->stream, s->mma_ev, 0));
    }

    qsplit_kernel<<<std::min(1024, (int)(((size_t)nqp * dim + 255) / 256)), 256, 0, c->stream>>>(sp.q, qmap, g, nqp, dim, qs);
    CU_OK(cudaGetLastError());
    CU_OK(cudaMemsetAsync(cand_cnt, 0, (size_t)2 * kMmaMaxNQ * 4, c->stream));   // candidate counts and grid-wide bounds

    if (get_encoder()) return -2;
    // CTA pairs (cta_group::2): each CTA of a pair stages half of the query tile, so the box is half as tall
    const bool pair = terms == 1 && tuning().mma_pair && (grid % 2) == 0 && grid >= 2;
    CUtensorMap tmap_q;
    cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)(2 * nqp)};
    cuuint64_t gstride[1] = {(cuuint64_t)dim * 4};
    const int bc = mma_box_cols(terms);
    cuuint32_t box[2] = {(cuuint32_t)bc, (cuuint32_t)(pair ? nqp / 2 : nqp)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(&tmap_q, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, qs, gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, bc == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(-2, "cuTensorMapEncodeTiled (queries) failed with CUresult %d", (int)r);

    MmaParams mp;
    memset(&mp, 0, sizeof(mp));
    mp.n_rows = sp.n_rows; mp.dim = dim; mp.n_chunks = (dim + bc - 1) / bc;
    mp.box_cols = bc;
    const CUtensorMap& tmap_keys = bc == 32 ? s->tmap : s->tmap_mma;   // the exact scan's [32 x 128] / 128-B-swizzle map serves here too
    mp.nq = g; mp.nqp = nqp; mp.f64 = norm ? 0 : 1;
    mp.knorm = sp.knorm; mp.lbest2 = lbest2; mp.gbound = gbound;
    mp.cand = c->mma_cand.as<uint64_t>(); mp.cand_cnt = cand_cnt;
    // rows that must reach the bound: the k-th best for the single-TF32 filter (its error band hangs below
    // that score), the (k + margin)-th for 3xTF32 (exactly that many are re-scored); two rows per CTA
    const int need = terms == 1 ? sp.topk : sp.topk + kMmaMargin;
    mp.r_need = (need + 1) / 2;
    mp.refresh_ns = 5000;
    mp.eps = terms == 1 ? eps : nullptr;
    mp.debug = tuning().mma_debug;
    mp.n_stages = mma_stage_count(nqp, terms, (size_t)s->max_smem, pair);
    if (mp.n_stages < 2) return fail(-4, "internal: tensor-core stages do not fit shared memory");
    const size_t smem = mma_smem_bytes(nqp, terms, mp.n_stages, pair);
    if (tuning().mma_trace)
        fprintf(stderr, "tfp mma: terms=%d pair=%d nq=%d nqp=%d stages=%d smem=%zu grid=%d\n", terms, (int)pair, g, nqp, mp.n_stages, smem, grid);
    auto launch_pass = [&]() -> int {
        if (pair) {
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(kMmaThreads); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            CU_OK((cudaLaunchKernelEx(&cfg, mma_scan_kernel<1, true>, tmap_keys, tmap_q, mp)));
        } else if (terms == 3) {
            mma_scan_kernel<3, false><<<grid, kMmaThreads, smem, c->stream>>>(tmap_keys, tmap_q, mp);
        } else {
            mma_scan_kernel<1, false><<<grid, kMmaThreads, smem, c->stream>>>(tmap_keys, tmap_q, mp);
        }
        CU_OK(cudaGetLastError());
        return 0;
    };
    mp.mode = 0;   // bound-only pass: one tile per CTA, so that every CTA owns a published best per query
    if (launch_pass()) return -1;
    mma_bound_kernel<<<(g + 7) / 8, 256, 0, c->stream>>>(lbest2, grid, mp.r_need, g, gbound);
    CU_OK(cudaGetLastError());
    mp.mode = 1;
    if (launch_pass()) return -1;
    if (serial) {
        std::lock_guard<std::mutex> g(s->mma_mu);
        CU_OK(cudaEventRecord(s->mma_ev, c->stream));
        s->mma_last = c;
    }

    FinalizeParams fp;
    memset(&fp, 0, sizeof(fp));
    fp.keys = sp.keys; fp.ids = sp.ids; fp.knorm = sp.knorm; fp.n_rows = sp.n_rows; fp.dim = dim;
    fp.q = sp.q; fp.qinfo = sp.qinfo; fp.q_first = q_base; fp.qmap = qmap;
    fp.cand = mp.cand; fp.cand_cnt = cand_cnt; fp.topk = sp.topk; fp.out_first = o_base;
    fp.out_cand = sp.out_cand; fp.out_ids = sp.out_ids; fp.out_sims = sp.out_sims; fp.out_slots = sp.out_slots;
    fp.out_n = sp.out_n; fp.out_rec = sp.out_rec; fp.rec_bytes = sp.rec_bytes; fp.rec_topk = sp.rec_topk;
    fp.status = status;
    fp.eps = eps;
    if (terms == 3) {
        if (norm) mma_finalize_kernel<false, 3><<<g, kFinThreads, kFinDynSmem, c->stream>>>(fp);
        else      mma_finalize_kernel<true, 3><<<g, kFinThreads, kFinDynSmem, c->stream>>>(fp);
    } else {
        if (norm) mma_finalize_kernel<false, 1><<<g, kFinThreads, kFinDynSmem, c->stream>>>(fp);
        else      mma_finalize_kernel<true, 1><<<g, kFinThreads, kFinDynSmem, c->stream>>>(fp);
    }
    CU_OK(cudaGetLastError());
    c->launches += 5;
    c->mma_passes++;
    CU_OK(cudaMemcpyAsync(c->h_status.p, status, (size_t)g * 4, cudaMemcpyDeviceToHost, c->stream));
    return 0;
}

// Tensor-core path for one group of g same-path queries (staged [q0, q0+g), outputs from slot o0).  The filter only
// has to be good enough to leave few rows to re-score exactly; when it is not -- scores so close together that the
// single-TF32 error band holds more rows than can be re-scored, or a candidate list that overflowed -- the flagged
// queries go, TOGETHER, through one more pass with the 3xTF32 filter (error ~1e-6 instead of ~2e-3), and only what
// that cannot settle either is re-run on the exact scan kernel, up to 8 queries per pass.
// The status words decide the escalation, so somebody has to wait for the GPU before the answer is final:
// mma_group_begin enqueues the first pass and the copy of its status words and returns; ctx_resolve_locked waits for
// them and enqueues whatever else the group needs.  The synchronous callers run one after the other; a caller-made
// context's tfp_find_ctx_launch leaves the second to the next call on the context (upload, launch, gather, download,
// sync, device results / records -- each starts with ctx_resolve_locked), so that launch returns without waiting.
struct MmaPending {
    LaunchPlan L;
    std::vector<int> idx;     // staged queries of the pass in flight, in status-word order
    int q0 = 0, o0 = 0, terms = 1;
    bool norm = false;
};

static int mma_group_begin(tfp_find_ctx* c, const LaunchPlan& L, int q0, int o0, int g, bool norm) {
    MmaPending* m = new MmaPending();
    m->L = L;
    m->idx.resize((size_t)g);
    for (int j = 0; j < g; j++) m->idx[(size_t)j] = q0 + j;
    m->q0 = q0; m->o0 = o0; m->norm = norm;
    m->terms = tuning().mma_terms;
    int rc = mma_pass_launch(c, m->L.p, m->idx.data(), g, q0, o0, norm, L.grid, m->terms);
    if (rc) { delete m; return rc; }
    c->pend = m;
    return 0;
}

static void ctx_drop_pending(tfp_find_ctx* c) {
    delete c->pend;
    c->pend = nullptr;
}

// Finish what mma_group_begin started on this context, if anything.  Caller holds the store's lock (shared or
// exclusive) and has made the store's device current.  Waits for the context's stream.
static int ctx_resolve_locked(tfp_find_ctx* c) {
    if (!c->pend) return 0;
    std::unique_ptr<MmaPending> m(c->pend);
    c->pend = nullptr;
    for (;;) {
        CU_OK(cudaStreamSynchronize(c->stream));
        if (tuning().mma_debug) return 0;   // (development timing runs skip the re-runs)
        const int* hs = c->h_status.as<int>();
        std::vector<int> flagged;
        for (size_t j = 0; j < m->idx.size(); j++) if (hs[j] != 0) flagged.push_back(m->idx[j]);
        if (flagged.empty()) return 0;
        if (m->terms == 1 && flagged.size() >= 3) {   // (one or two queries: an exact scan each costs no more than another pass)
            c->mma_escalated += (int64_t)flagged.size();
            m->idx.swap(flagged);
            m->terms = 3;
            int rc = mma_pass_launch(c, m->L.p, m->idx.data(), (int)m->idx.size(), m->q0, m->o0, m->norm, m->L.grid, 3);
            if (rc) return rc;
            continue;
        }
        // what neither filter settles is re-run on the exact kernel, up to 8 queries per pass over the keys
        c->mma_fallbacks += (int64_t)flagged.size();
        const int n_sel = (int)flagged.size();
        int done = 0;
        while (done < n_sel) {
            const int left = n_sel - done;
            int gg = left >= 8 ? 8 : (left >= 4 ? 4 : (left >= 2 ? 2 : 1)), launched = 0;
            int rc = exact_launch(c, m->L, m->q0, m->o0, gg, m->norm, &launched, flagged.data() + done);
            if (rc) return rc;
            done += launched;
        }
        return 0;
    }
}

// Enqueue the scan(s) for staged queries [first, first+count).  Caller holds the shared lock.
// xchg: push the finished records to the peer ranks from inside the scan kernel when the find is
// made of exact fused scans only (c->x_pushed tells the caller whether that happened).
// want_keys: the caller will ask for the key rows; exact fused scans then write them from their last CTA
// (c->keys_gathered), other paths leave them to gather_rows_kernel.
// local_records: the find never leaves this shard (tfp_store_find[_batch]): size the packed records by
// min(topk, rows) instead of the caller's topk -- only shards that exchange records need them equal.
// may_defer: see mma_group_begin.
static int ctx_launch_locked(tfp_find_ctx* c, int first, int count, int topk, const XchgParams* xchg = nullptr,
                             bool want_keys = false, bool local_records = false, bool may_defer = false) {
    tfp_store* s = c->st;
    if (int rc0 = ctx_resolve_locked(c)) return rc0;   // the previous launch's answer is finished before its buffers are reused
    c->x_pushed = false;
    c->keys_gathered = false;
    if (c->dim != s->dim)                                                            // store.go:480-482
        return fail(1, "Try to find key with length %d when existing length is %d", c->dim, s->dim);
    if (topk < 1) return fail(1, "opts.TopK = %d, must be >= 1", topk);             // store.go:484-486
    const int topk_host = topk;                              // the caller's arrays are strided by what the caller asked for
    if ((int64_t)topk > (int64_t)1 << 26) topk = 1 << 26;   // more than any shard can hold; keeps record sizes sane
    if (first < 0 || count <= 0 || first + count > c->nq) return fail(2, "find: query range [%d,%d) outside the %d staged queries", first, first + count, c->nq);
    const int64_t n = s->n;
    const int ksel = (int)std::min<int64_t>((int64_t)topk, std::max<int64_t>(n, 1));
    if (local_records) topk = ksel;
    c->last_count = count;
    c->last_topk = topk;
    c->last_topk_host = topk_host;
    c->last_stride = ksel;
    c->last_path = (s->normalized && c->qinfo[(size_t)first].normalized) ? TFP_PATH_NORMALIZED : TFP_PATH_FALLBACK;
    size_t slots = (size_t)count * ksel;
    const size_t rec_bytes = record_bytes(topk);
    {
        auto up16 = [](size_t v) { return (v + 15) & ~(size_t)15; };
        const size_t kd = (size_t)(s->dim > 0 ? s->dim : 1);
        c->off_ids = up16((size_t)count * 4);
        c->off_sims = c->off_ids + slots * 8;
        c->off_keys = up16(c->off_sims + slots * 4);
        const size_t off_cand = up16(c->off_keys + slots * kd * 4);
        const size_t off_slots = off_cand + slots * 8;
        const size_t off_rec = up16(off_slots + slots * 4);
        if (c->out_block.reserve(off_rec + (size_t)count * rec_bytes)) return -1;
        uint8_t* b = c->out_block.as<uint8_t>();
        c->out_n = (int*)b; c->out_ids = (int64_t*)(b + c->off_ids); c->out_sims = (float*)(b + c->off_sims);
        c->out_keys = (float*)(b + c->off_keys); c->out_cand = (uint64_t*)(b + off_cand);
        c->out_slots = (uint32_t*)(b + off_slots); c->out_rec = b + off_rec;
    }
    if (n == 0) {   // written-then-emptied store: empty result
        CU_OK(cudaMemsetAsync(c->out_n, 0, (size_t)count * 4, c->stream));
        CU_OK(cudaMemsetAsync(c->out_rec, 0, (size_t)count * rec_bytes, c->stream));
        return 0;
    }
    const int dim = s->dim;
    const bool fused = ksel <= TFP_FUSED_TOPK_MAX;
    const int n_chunks = (dim + kBoxCols - 1) / kBoxCols;
    const int dpad = n_chunks * kBoxCols;
    const int cap = fused ? std::max(512, next_pow2(ksel + std::max(ksel, 2 * kRowsPerTile))) : 2;
    const uint32_t n_tiles = (uint32_t)((n + kRowsPerTile - 1) / kRowsPerTile);
    // One persistent CTA per SM for the TMA scan (its stage ring keeps the loads in flight).  The generic kernel keeps
    // them in flight with resident warps instead: kGenCtasPerSm CTAs of four warps per SM.
    const int grid = (int)std::min<uint32_t>((uint32_t)std::min(s->sm_count * (s->tmap_ok ? 1 : kGenCtasPerSm), kMaxGrid), n_tiles);

    LaunchPlan L;
    ScanParams& p = L.p;
    memset(&p, 0, sizeof(p));
    L.fused = fused; L.dpad = dpad; L.cap = cap; L.grid = grid; L.ksel = ksel; L.count = count; L.first = first;
    L.n_tiles = n_tiles; L.n = n;
    p.keys = s->d_keys; p.ids = s->d_ids; p.knorm = s->d_knorm;
    p.n_rows = (uint32_t)n; p.dim = dim; p.n_chunks = n_chunks;
    p.topk = ksel; p.cap = cap;
    p.evict_first = ((size_t)n * dim * 4 > ((size_t)96 << 20)) ? 1 : 0;   // larger than L2: stream it
    p.q = c->d_q; p.qinfo = c->d_qinfo;
    p.gtau = c->sync_words.as<unsigned long long>();
    p.ticket = (unsigned int*)(c->sync_words.as<unsigned long long>() + kMaxNQ);
    p.tile_counter = (unsigned int*)(c->sync_words.as<unsigned long long>() + kMaxNQ + 1);
    p.lbest = c->sync_words.as<unsigned long long>() + kMaxNQ + 2;
    p.lbest_stride = kMaxGrid;
    p.m_local = (ksel + grid - 1) / grid;
    p.r_local = (ksel + p.m_local - 1) / p.m_local;
    p.refresh_ns = tuning().refresh_ns;
    if (tuning().disable_lbest) p.lbest = nullptr;
    if (tuning().scan_times) {   // development: per-CTA time stamps of the next scan, printed by tfp_find_ctx_sync
        if (c->dbg_times.reserve((size_t)(kMaxGrid * 4 + 8) * 8)) return -1;
        CU_OK(cudaMemsetAsync(c->dbg_times.p, 0, (size_t)(kMaxGrid * 4 + 8) * 8, c->stream));
        p.dbg_times = c->dbg_times.as<unsigned long long>();
        c->dbg_grid = grid;
    }
    p.out_cand = c->out_cand; p.out_ids = c->out_ids;
    p.out_sims = c->out_sims; p.out_slots = c->out_slots; p.out_n = c->out_n;
    p.out_rec = c->out_rec; p.rec_bytes = (int)rec_bytes; p.rec_topk = topk;

    if (fused) {
        if (c->part.reserve((size_t)grid * kMaxNQ * ksel * 8) || c->part_cnt.reserve((size_t)grid * kMaxNQ * 4)) return -1;
        p.part = c->part.as<uint64_t>(); p.part_cnt = c->part_cnt.as<int>();
    } else {
        if (c->all_cand.reserve((size_t)n * 8) || c->sorted_cand.reserve((size_t)n * 8)) return -1;
        p.all_cand = c->all_cand.as<uint64_t>();
    }

    auto launch_exact = [&](int qf, int of, int g, bool norm, int* launched, const int* sel = nullptr) -> int {
        return exact_launch(c, L, qf, of, g, norm, launched, sel);
    };
    auto path_of = [&](int qi) { return s->normalized && c->qinfo[(size_t)qi].normalized; };   // store.go:496

    // Tensor-core path: large same-path batches on a large store.  The lower bound it
    // relies on counts one row per CTA, so k + margin must not exceed the grid.
    const bool mma_ok = fused && s->tmap_ok && ksel <= kMmaMaxTopK && grid >= ksel + kMmaMargin &&
                        n >= tuning().mma_min_rows && tuning().mma_min_queries > 0 &&
                        mma_stage_count(kMmaMaxNQ, 3, (size_t)s->max_smem) >= 3;
    const bool all_exact_fused = fused && !(mma_ok && count >= tuning().mma_min_queries);
    if (xchg && all_exact_fused) {
        p.xchg = *xchg;     // every launch below is an exact fused scan: each pushes its own records
        c->x_pushed = true;
    }
    if (want_keys && all_exact_fused && (size_t)ksel == (size_t)c->last_stride) {
        p.out_keys = c->out_keys;   // ... and each copies its winners' rows
        c->keys_gathered = true;
    }
    int i = 0;
    while (i < count) {
        const bool norm = path_of(first + i);
        int run = 1;
        while (i + run < count && path_of(first + i + run) == norm) run++;
        if (mma_ok && run >= tuning().mma_min_queries) {
            const int g = std::min(run, kMmaMaxNQ);
            int rc = mma_group_begin(c, L, first + i, i, g, norm);
            if (rc) return rc;
            // a launch that is this one group and nothing else may leave the status words to the context's next call
            if (!(may_defer && i == 0 && g == count && tuning().mma_defer)) {
                rc = ctx_resolve_locked(c);
                if (rc) return rc;
            }
            i += g;
            continue;
        }
        int g = 1;
        if (fused) {
            g = std::min(run, kMaxNQ);
            g = g >= 8 ? 8 : (g >= 4 ? 4 : (g >= 2 ? 2 : 1));
        }
        int done = 0;
        int rc = launch_exact(first + i, i, g, norm, &done);
        if (rc) return rc;
        i += done;
    }
    return 0;
}

extern "C" int tfp_find_ctx_launch(tfp_find_ctx* c, int first, int count, int topk) {
    if (!c) return fail(2, "tfp_find_ctx_launch: ctx is NULL");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    return ctx_launch_locked(c, first, count, topk, nullptr, false, false, /*may_defer=*/true);
}

// Enqueue the gather of the winning rows of the last launch into c->out_keys.
static int ctx_gather_locked(tfp_find_ctx* c) {
    tfp_store* s = c->st;
    const int count = c->last_count, ks = c->last_stride, dim = s->dim;
    if (s->n == 0) return 0;
    dim3 grid((unsigned)ks, (unsigned)count);
    int threads = dim >= 256 ? 256 : (dim >= 64 ? 64 : 32);
    gather_rows_kernel<<<grid, threads, 0, c->stream>>>(s->d_keys, dim, c->out_slots, c->out_n, ks, c->out_keys);
    CU_OK(cudaGetLastError());
    c->launches++;
    return 0;
}

extern "C" int tfp_find_ctx_gather_keys(tfp_find_ctx* c, float** d_keys) {
    if (!c || !d_keys) return fail(2, "tfp_find_ctx_gather_keys: NULL argument");
    if (c->last_count <= 0) return fail(2, "tfp_find_ctx_gather_keys: nothing was launched on this context");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    int rc = ctx_resolve_locked(c);
    if (!rc) rc = ctx_gather_locked(c);
    if (rc) return rc;
    *d_keys = c->out_keys;
    return 0;
}

static int ctx_download_locked(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims, float* out_keys) {
    tfp_store* s = c->st;
    const int count = c->last_count, topk = c->last_topk_host, ks = c->last_stride, dim = s->dim;
    if (count <= 0) return fail(2, "find download: nothing was launched on this context");
    if (int rc = ctx_resolve_locked(c)) return rc;
    const size_t slots = (size_t)count * ks;
    // one D2H copy of the block's head: [n | ids | sims] and, when asked for, the key rows
    const size_t total = out_keys ? c->off_keys + slots * dim * 4 : c->off_sims + slots * 4;
    if (c->h_out.reserve(total)) return -1;
    uint8_t* h = (uint8_t*)c->h_out.p;
    if (s->n > 0) {
        if (out_keys && !c->keys_gathered && ctx_gather_locked(c)) return -1;
        CU_OK(cudaMemcpyAsync(h, c->out_block.p, total, cudaMemcpyDeviceToHost, c->stream));
    }
    CU_OK(cudaStreamSynchronize(c->stream));
    const int* hn = (const int*)h;
    for (int q = 0; q < count; q++) {
        int m = s->n > 0 ? hn[q] : 0;
        if (out_n) out_n[q] = m;
        if (out_ids) memcpy(out_ids + (size_t)q * topk, h + c->off_ids + (size_t)q * ks * 8, (size_t)m * 8);
        if (out_sims) memcpy(out_sims + (size_t)q * topk, h + c->off_sims + (size_t)q * ks * 4, (size_t)m * 4);
        if (out_keys) memcpy(out_keys + (size_t)q * topk * dim, h + c->off_keys + (size_t)q * ks * dim * 4, (size_t)m * dim * 4);
    }
    return 0;
}

extern "C" int tfp_find_ctx_download(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims, float* out_keys) {
    if (!c) return fail(2, "tfp_find_ctx_download: ctx is NULL");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    return ctx_download_locked(c, out_n, out_ids, out_sims, out_keys);
}
extern "C" int tfp_find_ctx_sync(tfp_find_ctx* c) {
    if (!c) return fail(2, "tfp_find_ctx_sync: ctx is NULL");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    if (int rc = ctx_resolve_locked(c)) return rc;
    CU_OK(cudaStreamSynchronize(c->stream));
    if (c->dbg_grid > 0 && c->dbg_times.p) {   // development: timeline of the last scan
        std::vector<unsigned long long> t((size_t)c->dbg_grid * 4);
        CU_OK(cudaMemcpy(t.data(), c->dbg_times.p, t.size() * 8, cudaMemcpyDeviceToHost));
        unsigned long long t0 = ~0ull, s_min = ~0ull, s_max = 0, e_max = 0, start_max = 0;
        double s_sum = 0;
        unsigned long long tl_min = ~0ull, tl_max = 0;
        for (int b = 0; b < c->dbg_grid; b++) {
            t0 = std::min(t0, t[b * 4]); start_max = std::max(start_max, t[b * 4]);
        }
        for (int b = 0; b < c->dbg_grid; b++) {
            s_min = std::min(s_min, t[b * 4 + 1]); s_max = std::max(s_max, t[b * 4 + 1]); s_sum += (double)(t[b * 4 + 1] - t0);
            e_max = std::max(e_max, t[b * 4 + 2]); tl_min = std::min(tl_min, t[b * 4 + 3]); tl_max = std::max(tl_max, t[b * 4 + 3]);
        }
        fprintf(stderr, "tfp scan timeline (us from first CTA start): last CTA start %.1f | scan end min %.1f avg %.1f max %.1f | "
                        "kernel end %.1f | tiles per CTA %llu..%llu\n",
                (start_max - t0) / 1e3, (s_min - t0) / 1e3, s_sum / c->dbg_grid / 1e3, (s_max - t0) / 1e3, (e_max - t0) / 1e3,
                tl_min, tl_max);
        unsigned long long ls[8];
        CU_OK(cudaMemcpy(ls, c->dbg_times.as<unsigned long long>() + (size_t)kMaxGrid * 4, sizeof(ls), cudaMemcpyDeviceToHost));
        if (ls[0])
            fprintf(stderr, "tfp last CTA (us from first CTA start): ticket %.1f | lists counted %.1f (%llu words) | merged %.1f | "
                            "sorted %.1f | written %.1f\n",
                    (ls[0] - t0) / 1e3, (ls[1] - t0) / 1e3, ls[5], (ls[2] - t0) / 1e3, (ls[3] - t0) / 1e3, (ls[4] - t0) / 1e3);
        c->dbg_grid = 0;
    }
    return 0;
}
extern "C" void* tfp_find_ctx_stream(tfp_find_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int64_t tfp_find_ctx_launch_count(tfp_find_ctx* c) { return c ? c->launches : 0; }
extern "C" int tfp_find_ctx_mma_counters(tfp_find_ctx* c, int64_t* passes, int64_t* exact_reruns) {
    if (!c) return fail(2, "tfp_find_ctx_mma_counters: ctx is NULL");
    if (passes) *passes = c->mma_passes;
    if (exact_reruns) *exact_reruns = c->mma_fallbacks;
    return 0;
}
extern "C" int64_t tfp_find_ctx_mma_escalated(tfp_find_ctx* c) { return c ? c->mma_escalated : 0; }
extern "C" int tfp_find_ctx_device_results(tfp_find_ctx* c, uint64_t** d_cand, int64_t** d_ids, float** d_sims, int** d_n) {
    if (!c) return fail(2, "tfp_find_ctx_device_results: ctx is NULL");
    if (c->pend) {   // the arrays are final once the tensor-core batch in flight has been looked at
        std::shared_lock<RwMutex> lk(c->st->rw);
        if (set_device(c->st) || ctx_resolve_locked(c)) return -1;
    }
    if (d_cand) *d_cand = c->out_cand;
    if (d_ids) *d_ids = c->out_ids;
    if (d_sims) *d_sims = c->out_sims;
    if (d_n) *d_n = c->out_n;
    return c->last_stride;   // entries per query in the device arrays (min(topk, rows))
}

// ---------------------------------------------------------------------------
// synchronous finds (what the Go host calls)
// ---------------------------------------------------------------------------
static int store_find_batch(tfp_store* s, const float* q, int nq, int dim, int topk, int* out_n, int64_t* out_ids,
                            float* out_sims, float* out_keys, int* out_path) {
    if (!s) return fail(2, "tfp_store_find: store is NULL");
    std::shared_lock<RwMutex> lk(s->rw);
    if (dim != s->dim)                                                               // store.go:480-482 (fires on -1 too)
        return fail(1, "Try to find key with length %d when existing length is %d", dim, s->dim);
    if (topk < 1) return fail(1, "opts.TopK = %d, must be >= 1", topk);             // store.go:484-486
    if (!q || nq <= 0) return fail(2, "tfp_store_find: q is NULL or nq <= 0");
    if (set_device(s)) return -1;
    std::lock_guard<std::mutex> g(s->find_mu);
    tfp_find_ctx* c = s->default_ctx;
    int rc = ctx_upload_locked(c, q, nq, dim);
    if (rc) return rc;
    rc = ctx_launch_locked(c, 0, nq, topk, nullptr, out_keys != nullptr, /*local_records=*/true);
    if (rc) return rc;
    if (out_path) *out_path = c->last_path;   // read while find_mu is held: another thread's find cannot slip in
    return ctx_download_locked(c, out_n, out_ids, out_sims, out_keys);
}

// ---------------------------------------------------------------------------
// cross-shard merge
// ---------------------------------------------------------------------------
extern "C" size_t tfp_record_bytes(int topk) { return record_bytes(topk < 0 ? 0 : topk); }

extern "C" int tfp_find_ctx_device_records(tfp_find_ctx* c, void** d_records, size_t* rec_bytes) {
    if (!c || !d_records || !rec_bytes) return fail(2, "tfp_find_ctx_device_records: NULL argument");
    if (c->last_count <= 0) return fail(2, "tfp_find_ctx_device_records: nothing was launched on this context");
    if (c->pend) {
        std::shared_lock<RwMutex> lk(c->st->rw);
        if (set_device(c->st)) return -1;
        if (int rc = ctx_resolve_locked(c)) return rc;
    }
    *d_records = c->out_rec;
    *rec_bytes = record_bytes(c->last_topk);
    return 0;
}

extern "C" int tfp_merge_records(const void* d_gathered, int n_shards, int nq, int topk, float* o_sims, int64_t* o_ids,
                                 int* o_shard, int* o_index, int* o_n, void* stream) {
    if (!d_gathered || !o_sims || !o_ids || !o_shard || !o_index || !o_n)
        return fail(2, "tfp_merge_records: NULL argument");
    if (n_shards < 1 || nq < 1 || topk < 1) return fail(2, "tfp_merge_records: bad sizes");
    int64_t total = (int64_t)n_shards * topk;
    if (total > 16384) return fail(2, "tfp_merge_records: n_shards x topk = %lld exceeds 16384", (long long)total);
    int P = next_pow2((int)std::max<int64_t>(total, 2));
    if ((size_t)P * 8 > 48 * 1024)
        CU_OK(cudaFuncSetAttribute(merge_records_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P * 8));
    merge_records_kernel<<<nq, 256, (size_t)P * 8, (cudaStream_t)stream>>>((const uint8_t*)d_gathered,
                                                                           (int)record_bytes(topk),
                                                                           record_bytes(topk) * (size_t)nq, n_shards, nq, topk, P,
                                                                           o_sims, o_ids, o_shard, o_index, o_n, nullptr, 0, 0u);
    CU_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// tfp_scan_topk_noalloc: the same scan + fused top-k, with every byte of scratch in a
// fixed array that is part of the library image (see include/tfp_store.h for who needs
// this and why).  Nothing below may call cudaMalloc, cudaFree or anything that allocates.
// ---------------------------------------------------------------------------
constexpr int kNoAllocMaxGrid = 160;   // CTAs per launch are capped so the candidate lists fit the fixed array
struct NoAllocLayout {                 // byte offsets inside g_noalloc_scratch
    static constexpr size_t part = 0;                                                        // [grid x kMaxNQ x topk] u64
    static constexpr size_t part_cnt = part + (size_t)kNoAllocMaxGrid * kMaxNQ * TFP_NOALLOC_TOPK_MAX * 8;
    static constexpr size_t sync_words = part_cnt + (size_t)kNoAllocMaxGrid * kMaxNQ * 4;    // gtau[kMaxNQ] + ticket + tile counter
    static constexpr size_t out_cand = sync_words + (size_t)(kMaxNQ + 2 + kMaxNQ * kNoAllocMaxGrid) * 8;
    static constexpr size_t out_slots = out_cand + (size_t)TFP_NOALLOC_TOPK_MAX * 8;
    static constexpr size_t out_rec = out_slots + (size_t)TFP_NOALLOC_TOPK_MAX * 4;
    static constexpr size_t qinfo = out_rec + 1024;
    static constexpr size_t end = qinfo + 64;
};
static_assert(NoAllocLayout::end <= TFP_NOALLOC_SCRATCH_BYTES, "the fixed scratch array is too small for its layout");

// In host code this name is only a handle for cudaGetSymbolAddress; it is NOT a device address.
static __device__ __align__(256) unsigned char g_noalloc_scratch[TFP_NOALLOC_SCRATCH_BYTES];

struct NoAllocDevice {
    bool ready = false;
    unsigned char* scratch = nullptr;   // device address of g_noalloc_scratch on this device
    int sm_count = 0;
    int max_smem = 0;
};
static NoAllocDevice g_noalloc_dev[64];
static std::mutex g_noalloc_mu;

// Once per device, with that device current: resolve the array's device address, read the
// device limits, opt the kernels into large dynamic shared memory.  None of it allocates.
static int noalloc_setup(int dev, NoAllocDevice** out) {
    std::lock_guard<std::mutex> g(g_noalloc_mu);
    NoAllocDevice& d = g_noalloc_dev[dev];
    if (!d.ready) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail(-1, "cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
        int cc_major = 0;
        CU_OK(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev));
        if (cc_major != 10) return fail(-1, "device %d is not sm_100; libtfpstore is built for sm_100a (B200) only", dev);
        CU_OK(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
        CU_OK(cudaDeviceGetAttribute(&d.max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        CU_OK((optin_tma<1, false, false>(d.max_smem)));
        CU_OK((optin_gen<1, false, false>(d.max_smem)));
        d.scratch = (unsigned char*)ptr;
        d.ready = true;
    }
    *out = &d;
    return 0;
}

extern "C" int tfp_scan_topk_noalloc(const float* d_keys, int64_t n_rows, int dim, const int64_t* d_ids,
                                     const float* d_query, int topk, float* d_out_sims, int64_t* d_out_ids,
                                     int* d_out_n, void* stream) {
    if (!d_keys || !d_ids || !d_query || !d_out_sims || !d_out_ids || !d_out_n)
        return fail(2, "tfp_scan_topk_noalloc: NULL argument");
    if (topk < 1) return fail(1, "opts.TopK = %d, must be >= 1", topk);
    if (topk > TFP_NOALLOC_TOPK_MAX)
        return fail(2, "tfp_scan_topk_noalloc: topk = %d exceeds TFP_NOALLOC_TOPK_MAX = %d (the scratch array is fixed)",
                    topk, TFP_NOALLOC_TOPK_MAX);
    if (dim < 1 || dim > TFP_NOALLOC_DIM_MAX)
        return fail(2, "tfp_scan_topk_noalloc: dim = %d outside [1, TFP_NOALLOC_DIM_MAX = %d]", dim, TFP_NOALLOC_DIM_MAX);
    if (n_rows < 0 || n_rows >= 0xFFFFFFF0ll)
        return fail(2, "tfp_scan_topk_noalloc: n_rows = %lld outside [0, 2^32 - 16)", (long long)n_rows);
    int dev = 0;
    CU_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(2, "tfp_scan_topk_noalloc: device ordinal %d not supported", dev);
    NoAllocDevice* nd = nullptr;
    if (noalloc_setup(dev, &nd)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rows == 0) {
        CU_OK(cudaMemsetAsync(d_out_n, 0, sizeof(int), st));
        return 0;
    }
    unsigned char* base = nd->scratch;   // the address from cudaGetSymbolAddress, never the host-side symbol
    const int ksel = (int)std::min<int64_t>((int64_t)topk, n_rows);
    const int n_chunks = (dim + kBoxCols - 1) / kBoxCols;
    const int dpad = n_chunks * kBoxCols;
    const int cap = std::max(512, next_pow2(ksel + std::max(ksel, 2 * kRowsPerTile)));
    const uint32_t n_tiles = (uint32_t)((n_rows + kRowsPerTile - 1) / kRowsPerTile);
    const int grid = (int)std::min<uint32_t>((uint32_t)std::min(nd->sm_count, kNoAllocMaxGrid), n_tiles);

    ScanParams p;
    memset(&p, 0, sizeof(p));
    p.keys = d_keys; p.ids = d_ids; p.knorm = nullptr;
    p.n_rows = (uint32_t)n_rows; p.dim = dim; p.n_chunks = n_chunks;
    p.topk = ksel; p.cap = cap;
    p.evict_first = ((size_t)n_rows * dim * 4 > ((size_t)96 << 20)) ? 1 : 0;
    p.q = d_query; p.qinfo = (const QInfo*)(base + NoAllocLayout::qinfo);   // not read on the fp32 path
    p.q_first = 0; p.out_first = 0;
    p.part = (uint64_t*)(base + NoAllocLayout::part);
    p.part_cnt = (int*)(base + NoAllocLayout::part_cnt);
    p.gtau = (unsigned long long*)(base + NoAllocLayout::sync_words);
    p.ticket = (unsigned int*)(base + NoAllocLayout::sync_words + kMaxNQ * 8);
    p.tile_counter = (unsigned int*)(base + NoAllocLayout::sync_words + (kMaxNQ + 1) * 8);
    p.lbest = (unsigned long long*)(base + NoAllocLayout::sync_words + (kMaxNQ + 2) * 8);
    p.lbest_stride = kNoAllocMaxGrid;
    p.m_local = (ksel + grid - 1) / grid;
    p.r_local = (ksel + p.m_local - 1) / p.m_local;
    p.refresh_ns = tuning().refresh_ns;
    if (tuning().disable_lbest) p.lbest = nullptr;
    p.out_cand = (uint64_t*)(base + NoAllocLayout::out_cand);
    p.out_slots = (uint32_t*)(base + NoAllocLayout::out_slots);
    p.out_rec = base + NoAllocLayout::out_rec;
    p.rec_bytes = (int)record_bytes(ksel); p.rec_topk = ksel;
    p.out_ids = d_out_ids; p.out_sims = d_out_sims; p.out_n = d_out_n;

    // The array keeps whatever the previous call left.  The grid-wide bounds (gtau, lbest), the
    // ticket and the tile counter must start at zero; everything else is written before it is read.
    CU_OK(cudaMemsetAsync(base + NoAllocLayout::sync_words, 0, (kMaxNQ + 2 + kMaxNQ * kNoAllocMaxGrid) * 8, st));

    // TMA needs 16-byte aligned rows and a box that fits the tensor; otherwise the generic kernel.
    bool use_tma = dim >= kBoxCols && (dim % 4) == 0 && n_rows >= kRowsPerTile &&
                   (((uintptr_t)d_keys | (uintptr_t)d_query) & 15) == 0;
    int n_stages = 0;
    size_t smem = 0;
    if (use_tma) {
        size_t fixed = scan_smem_bytes(0, 1, dpad, cap) + 2 * kMaxStages * 8;
        n_stages = std::min(kMaxStages, ((int)nd->max_smem - (int)fixed) / kStageBytes);
        if (n_stages >= 3) smem = scan_smem_bytes(n_stages, 1, dpad, cap);
        else use_tma = false;
    }
    if (use_tma) {
        if (get_encoder()) return -2;
        CUtensorMap tmap;
        cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)n_rows};
        cuuint64_t gstride[1] = {(cuuint64_t)dim * 4};
        cuuint32_t box[2] = {(cuuint32_t)kBoxCols, (cuuint32_t)kRowsPerTile};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = g_encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)d_keys, gdim, gstride, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(-2, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        p.n_stages = n_stages;
        scan_tma_kernel<1, false, false><<<grid, kScanThreads, smem, st>>>(tmap, p);
    } else {
        smem = scan_generic_smem_bytes(1, cap);
        scan_generic_kernel<1, false, false><<<grid, kSelThreads, smem, st>>>(p);
    }
    CU_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// sharded find in one call: scan + ncclAllGather(records) + merge on the context's stream
// ---------------------------------------------------------------------------
static int nccl_api() {
    std::call_once(g_nccl_once, [] {
        // prefer the copy the process already loaded (the host's own NCCL), then the system one
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
        g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
        g_nccl.AllGather = (decltype(g_nccl.AllGather))dlsym(h, "ncclAllGather");
        g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
        g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
        g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.AllGather && g_nccl.CommDestroy;
    });
    return g_nccl.ok ? 0 : fail(-5, "NCCL (libnccl.so.2) could not be resolved in this process");
}
#define NCCL_OK(expr)                                                                                    \
    do {                                                                                                 \
        int r__ = (expr);                                                                                \
        if (r__ != 0)                                                                                    \
            return fail(-5, "NCCL error %d (%s) at %s:%d", r__,                                          \
                        g_nccl.GetErrorString ? g_nccl.GetErrorString(r__) : "?", __FILE__, __LINE__);   \
    } while (0)

extern "C" int tfp_nccl_unique_id(void* out_id, size_t out_bytes) {
    if (!out_id || out_bytes < sizeof(NcclId)) return fail(2, "tfp_nccl_unique_id: need a 128-byte buffer");
    if (nccl_api()) return -5;
    NCCL_OK(g_nccl.GetUniqueId(out_id));
    return 0;
}

extern "C" int tfp_find_ctx_comm_init(tfp_find_ctx* c, const void* unique_id, int rank, int world) {
    if (!c || !unique_id || world < 1 || rank < 0 || rank >= world) return fail(2, "tfp_find_ctx_comm_init: bad argument");
    if (nccl_api()) return -5;
    if (set_device(c->st)) return -1;
    if (c->nccl_comm && c->comm_owned) g_nccl.CommDestroy(c->nccl_comm);
    c->nccl_comm = nullptr;
    NcclId id;
    memcpy(&id, unique_id, sizeof(id));
    NCCL_OK(g_nccl.CommInitRank(&c->nccl_comm, world, id, rank));
    c->comm_owned = true;
    c->comm_rank = rank;
    c->comm_world = world;
    return 0;
}

extern "C" int tfp_find_ctx_comm_share(tfp_find_ctx* c, tfp_find_ctx* owner) {
    if (!c || !owner || !owner->nccl_comm || c->st != owner->st) return fail(2, "tfp_find_ctx_comm_share: bad argument");
    if (c->nccl_comm && c->comm_owned && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl_comm);
    c->nccl_comm = owner->nccl_comm;
    c->comm_owned = false;
    c->comm_rank = owner->comm_rank;
    c->comm_world = owner->comm_world;
    return 0;
}

// ---------------------------------------------------------------------------
// Peer exchange for sharded finds: the scan kernel itself writes this shard's records into every
// rank's gather block over NVLink (cudaIpc-mapped peer memory) and raises a sequence word there; the
// receiving stream waits on its words with a stream memory operation, then merges.  No collective
// library call and no extra kernel per find.  Buffer reuse is safe with two parities: a rank pushes
// find s+2 only after its own wait for find s+1, and a peer's word s+1 is raised by its scan s+1,
// which its stream runs after its merge of find s -- the last reader of the parity being overwritten.
// ---------------------------------------------------------------------------
typedef CUresult (*PFN_streamWaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static PFN_streamWaitValue32 g_wait32 = nullptr;
static std::once_flag g_wait32_once;
static int wait32_api() {
    std::call_once(g_wait32_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            g_wait32 = (PFN_streamWaitValue32)fn;
    });
    return g_wait32 ? 0 : fail(-7, "cuStreamWaitValue32 is not available from the CUDA driver");
}

static size_t xbuf_flags_offset(int world, size_t slot) { return ((size_t)2 * world * slot + 255) & ~(size_t)255; }

static void p2p_detach(tfp_find_ctx* c) {
    for (int r = 0; r < kXchgMaxWorld; r++) {
        if (c->x_peer[r] && r != c->x_me && !c->x_root_mode) cudaIpcCloseMemHandle(c->x_peer[r]);
        c->x_peer[r] = nullptr;
    }
    c->x_ready = false;
}

extern "C" int tfp_find_ctx_p2p_export(tfp_find_ctx* c, int world, size_t slot_bytes, void* handle_out, size_t handle_bytes) {
    if (!c || !handle_out || handle_bytes < sizeof(cudaIpcMemHandle_t)) return fail(2, "tfp_find_ctx_p2p_export: bad argument");
    if (world < 2 || world > kXchgMaxWorld) return fail(2, "tfp_find_ctx_p2p_export: world = %d outside [2, %d]", world, kXchgMaxWorld);
    if (slot_bytes < 256 || slot_bytes > ((size_t)64 << 20)) return fail(2, "tfp_find_ctx_p2p_export: slot_bytes outside [256, 64 MiB]");
    std::unique_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st) || wait32_api()) return -1;
    CU_OK(cudaStreamSynchronize(c->stream));
    p2p_detach(c);
    slot_bytes = (slot_bytes + 255) & ~(size_t)255;
    const size_t total = xbuf_flags_offset(world, slot_bytes) + 256;
    c->xbuf.release();                      // a fresh allocation: an exported block is never resized in place
    if (c->xbuf.reserve(total)) return -1;
    CU_OK(cudaMemset(c->xbuf.p, 0, total));  // the sequence words start at 0, before anybody can see the handle
    c->x_slot = slot_bytes;
    c->x_world = world;
    c->x_seq = 0;
    cudaIpcMemHandle_t h;
    CU_OK(cudaIpcGetMemHandle(&h, c->xbuf.p));
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}

extern "C" int tfp_find_ctx_p2p_attach(tfp_find_ctx* c, const void* handles, int rank, int world) {
    if (!c || !handles) return fail(2, "tfp_find_ctx_p2p_attach: NULL argument");
    if (!c->xbuf.p || world != c->x_world || rank < 0 || rank >= world)
        return fail(2, "tfp_find_ctx_p2p_attach: call tfp_find_ctx_p2p_export with the same world first");
    std::unique_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st) || wait32_api()) return -1;
    c->x_me = rank;
    for (int r = 0; r < world; r++) {
        if (r == rank) { c->x_peer[r] = c->xbuf.as<uint8_t>(); continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const uint8_t*)handles + (size_t)r * sizeof(h), sizeof(h));
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            p2p_detach(c);
            return fail(-7, "cudaIpcOpenMemHandle for rank %d failed: %s (no peer access between these GPUs?)", r, cudaGetErrorString(e));
        }
        c->x_peer[r] = (uint8_t*)ptr;
    }
    // prove that this stream accepts the wait operation before any find depends on it (0 >= 0: passes at once)
    unsigned int* words = (unsigned int*)(c->xbuf.as<uint8_t>() + xbuf_flags_offset(world, c->x_slot));
    CUresult r = g_wait32((CUstream)c->stream, (CUdeviceptr)(words + rank), 0u, CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) { p2p_detach(c); return fail(-7, "cuStreamWaitValue32 failed with CUresult %d", (int)r); }
    CU_OK(cudaStreamSynchronize(c->stream));
    c->x_ready = true;
    return 0;
}

extern "C" int tfp_find_ctx_p2p_detach(tfp_find_ctx* c) {
    if (!c) return fail(2, "tfp_find_ctx_p2p_detach: ctx is NULL");
    std::unique_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    CU_OK(cudaStreamSynchronize(c->stream));
    p2p_detach(c);
    return 0;
}

struct MergedLayout {   // [count x topk] sims f32 | ids i64 | shard i32 | index i32 | n[count] i32 (each 16-aligned)
    size_t sims, ids, shard, index, n, total;
    MergedLayout(int k, int count) {
        auto up = [](size_t v) { return (v + 15) & ~(size_t)15; };
        const size_t e = (size_t)k * count;
        sims = 0; ids = up(e * 4); shard = ids + e * 8; index = up(shard + e * 4);
        n = up(index + e * 4); total = up(n + (size_t)count * 4);
    }
};

// wait_words: the peers' sequence words the merge kernel itself waits on (peer exchange), or nullptr.
static int launch_merge(tfp_find_ctx* c, const uint8_t* gathered, size_t shard_stride, size_t rb, int world, int count, int k,
                        const unsigned int* wait_words = nullptr, int wait_me = 0, unsigned int wait_seq = 0) {
    MergedLayout L(k, count);
    if (c->merged.reserve(L.total)) return -1;
    uint8_t* m = c->merged.as<uint8_t>();
    int P = next_pow2(std::max(world * k, 2));
    if ((size_t)P * 8 > 48 * 1024)
        CU_OK(cudaFuncSetAttribute(merge_records_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P * 8));
    merge_records_kernel<<<count, 256, (size_t)P * 8, c->stream>>>(gathered, (int)rb, shard_stride, world, count, k, P,
                                                                  (float*)(m + L.sims), (int64_t*)(m + L.ids),
                                                                  (int*)(m + L.shard), (int*)(m + L.index), (int*)(m + L.n),
                                                                  wait_words, wait_me, wait_seq);
    CU_OK(cudaGetLastError());
    c->launches++;
    c->merged_topk = k;
    c->merged_count = count;
    return 0;
}

// Peer exchange of one sharded find (see the block comment above tfp_find_ctx_p2p_export).
// with_keys: the winning key rows travel too, so that the merged answer carries StoresFindResult.Keys.
static int p2p_merge_locked(tfp_find_ctx* c);

// defer_merge: enqueue the scan and the pushes only; the caller enqueues the merge later with p2p_merge_locked
// (one process over several GPUs: shard 0's waiting merge kernel is launched once every shard's scan is in its
// stream, so that nothing the host still has to do can end up waiting behind a kernel that waits for it).
static int p2p_launch_locked(tfp_find_ctx* c, int first, int count, int topk, bool with_keys, bool defer_merge = false) {
    tfp_store* s = c->st;
    const unsigned int seq = c->x_seq + 1;
    const size_t par_off = (size_t)(seq & 1u) * c->x_world * c->x_slot;
    const size_t flags_off = xbuf_flags_offset(c->x_world, c->x_slot);
    XchgParams x;
    memset(&x, 0, sizeof(x));
    x.world = c->x_world; x.me = c->x_me; x.seq = seq; x.raise = 1;
    for (int r = 0; r < c->x_world; r++) {
        x.slot[r] = c->x_peer[r] + par_off + (size_t)c->x_me * c->x_slot;
        x.flag[r] = (unsigned int*)(c->x_peer[r] + flags_off) + c->x_me;
    }
    if (c->x_root_mode) {
        // One process, several GPUs (multi_store.inl): records are gathered on shard 0 only.  The kernels see a
        // one-destination exchange: shard 0 writes its own slot and raises nothing (its merge follows in the same
        // stream), every other shard writes into shard 0's block over NVLink and raises its word there.
        x.world = 1;
        x.slot[0] = c->x_peer[0] + par_off + (size_t)c->x_me * c->x_slot;
        x.flag[0] = (unsigned int*)(c->x_peer[0] + flags_off) + c->x_me;
    }
    // records travel inside the scan kernel when nothing else has to be pushed before the words are raised
    int rc = ctx_launch_locked(c, first, count, topk, with_keys ? nullptr : &x);
    if (rc) return rc;     // argument errors: nothing was enqueued and seq was not consumed (same on every rank)
    c->x_seq = seq;
    const int k = c->last_topk, dim = s->dim;
    const size_t rb = record_bytes(k);
    const size_t keys_off = (rb * count + 255) & ~(size_t)255;
    if (with_keys && s->n > 0) {
        // winning rows go from the store straight into every rank's slot (stride topk, whatever this shard holds)
        dim3 grid((unsigned)c->last_stride, (unsigned)count);
        push_rows_kernel<<<grid, 256, 0, c->stream>>>(x, s->d_keys, dim, c->out_slots, c->out_n, c->last_stride, k, keys_off);
        CU_OK(cudaGetLastError());
        c->launches++;
    }
    if (!c->x_pushed) {
        push_records_kernel<<<x.world, 256, 0, c->stream>>>(x, c->out_rec, rb * count);
        CU_OK(cudaGetLastError());
        c->launches++;
    }
    c->x_m_count = count;
    c->x_m_with_keys = with_keys;
    if (c->x_root_mode && c->x_me != 0) return 0;   // only shard 0 merges
    if (defer_merge) return 0;
    return p2p_merge_locked(c);
}

// Second half of a peer-exchange find: wait for the peers' records of find x_seq, merge, gather the winners' rows.
static int p2p_merge_locked(tfp_find_ctx* c) {
    tfp_store* s = c->st;
    const unsigned int seq = c->x_seq;
    const int count = c->x_m_count, k = c->last_topk, dim = s->dim;
    const bool with_keys = c->x_m_with_keys;
    const size_t par_off = (size_t)(seq & 1u) * c->x_world * c->x_slot;
    const size_t flags_off = xbuf_flags_offset(c->x_world, c->x_slot);
    const size_t rb = record_bytes(k);
    const size_t keys_off = (rb * count + 255) & ~(size_t)255;
    unsigned int* words = (unsigned int*)(c->xbuf.as<uint8_t>() + flags_off);
    // The merge kernel waits for the peers' words itself (default).  TFP_XCHG_STREAM_WAIT=1 keeps the wait in the
    // stream instead: one cuStreamWaitValue32 per peer, measured 7 driver calls + a late kernel launch per find.
    const bool kernel_wait = !tuning().xchg_stream_wait && !c->x_no_spin;
    if (!kernel_wait && !c->x_no_spin) {
        if (wait32_api()) return -7;
        for (int r = 0; r < c->x_world; r++) {
            if (r == c->x_me) continue;
            CUresult wr = g_wait32((CUstream)c->stream, (CUdeviceptr)(words + r), seq, CU_STREAM_WAIT_VALUE_GEQ);
            if (wr != CUDA_SUCCESS) return fail(-7, "cuStreamWaitValue32 failed with CUresult %d", (int)wr);
        }
    }
    const uint8_t* block = c->xbuf.as<uint8_t>() + par_off;
    MergedLayout L(k, count);
    const size_t want = L.total + (with_keys ? (size_t)count * k * dim * 4 : 0);
    if (c->merged.reserve(want)) return -1;
    int rc = launch_merge(c, block, c->x_slot, rb, c->x_world, count, k, kernel_wait ? words : nullptr, c->x_me, seq);
    if (rc) return rc;
    c->merged_keys_dim = 0;
    if (with_keys) {
        uint8_t* m = c->merged.as<uint8_t>();
        dim3 grid((unsigned)k, (unsigned)count);
        gather_merged_rows_kernel<<<grid, 256, 0, c->stream>>>(block, c->x_slot, keys_off, dim, k, (const int*)(m + L.shard),
                                                              (const int*)(m + L.index), (const int*)(m + L.n),
                                                              (float*)(m + L.total));
        CU_OK(cudaGetLastError());
        c->launches++;
        c->merged_keys_dim = dim;
    }
    return 0;
}

static int sharded_check(tfp_find_ctx* c, int topk, int world) {
    // sizes depend on the arguments only, so every rank takes the same branch
    const int64_t kk = std::min<int64_t>(std::max(topk, 1), (int64_t)1 << 26);
    if ((int64_t)world * kk > 16384) return fail(2, "sharded find: world x topk = %lld exceeds 16384", (long long)world * kk);
    (void)c;
    return 0;
}

extern "C" int tfp_find_ctx_launch_sharded_batch(tfp_find_ctx* c, int first, int count, int topk) {
    if (!c) return fail(2, "tfp_find_ctx_launch_sharded: ctx is NULL");
    if (!c->nccl_comm && !c->x_ready)
        return fail(2, "tfp_find_ctx_launch_sharded: call tfp_find_ctx_p2p_attach or tfp_find_ctx_comm_init first");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    const int world = c->x_ready ? c->x_world : c->comm_world;
    if (sharded_check(c, topk, world)) return 2;
    const size_t rb_want = record_bytes(std::max(topk, 1));
    const bool p2p = c->x_ready && count > 0 && rb_want * (size_t)count <= c->x_slot;
    if (!p2p && !c->nccl_comm)
        return fail(2, "sharded find: %d records of %zu bytes exceed the peer-exchange slot (%zu bytes) and no NCCL communicator is set",
                    count, rb_want, c->x_slot);
    if (p2p) return p2p_launch_locked(c, first, count, topk, false);
    int rc = ctx_launch_locked(c, first, count, topk);
    if (rc) return rc;
    const int k = c->last_topk;
    const size_t rb = record_bytes(k);
    if (c->gathered.reserve(rb * count * world)) return -1;
    NCCL_OK(g_nccl.AllGather(c->out_rec, c->gathered.p, rb * count, /*ncclUint8*/ 1, c->nccl_comm, c->stream));
    c->merged_keys_dim = 0;
    return launch_merge(c, c->gathered.as<uint8_t>(), rb * count, rb, world, count, k);
}

extern "C" size_t tfp_p2p_slot_bytes_with_keys(int count, int topk, int dim) {
    if (count < 1 || topk < 1 || dim < 1) return 0;
    return ((record_bytes(topk) * (size_t)count + 255) & ~(size_t)255) + (size_t)count * topk * dim * 4;
}

extern "C" int tfp_find_ctx_launch_sharded_keys(tfp_find_ctx* c, int first, int count, int topk) {
    if (!c) return fail(2, "tfp_find_ctx_launch_sharded_keys: ctx is NULL");
    if (!c->x_ready) return fail(2, "tfp_find_ctx_launch_sharded_keys: call tfp_find_ctx_p2p_attach first");
    std::shared_lock<RwMutex> lk(c->st->rw);
    if (set_device(c->st)) return -1;
    if (sharded_check(c, topk, c->x_world)) return 2;
    if (count < 1 || c->st->dim < 1 || tfp_p2p_slot_bytes_with_keys(count, std::max(topk, 1), c->st->dim) > c->x_slot)
        return fail(2, "sharded find with keys: %d queries x top-%d x %d floats exceed the peer-exchange slot (%zu bytes)",
                    count, topk, c->st->dim, c->x_slot);
    return p2p_launch_locked(c, first, count, topk, true);
}

extern "C" int tfp_find_ctx_launch_sharded(tfp_find_ctx* c, int qi, int topk) {
    return tfp_find_ctx_launch_sharded_batch(c, qi, 1, topk);
}

extern "C" int tfp_find_ctx_merged_results(tfp_find_ctx* c, float** d_sims, int64_t** d_ids, int** d_shard, int** d_index,
                                           int** d_n) {
    if (!c || c->merged_topk <= 0) return fail(2, "tfp_find_ctx_merged_results: no sharded launch on this context");
    MergedLayout L(c->merged_topk, c->merged_count);
    uint8_t* m = c->merged.as<uint8_t>();
    if (d_sims) *d_sims = (float*)(m + L.sims);
    if (d_ids) *d_ids = (int64_t*)(m + L.ids);
    if (d_shard) *d_shard = (int*)(m + L.shard);
    if (d_index) *d_index = (int*)(m + L.index);
    if (d_n) *d_n = (int*)(m + L.n);
    return 0;
}

extern "C" int tfp_find_ctx_download_merged(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims, int* out_shard,
                                            int* out_index) {
    if (!c || c->merged_topk <= 0) return fail(2, "tfp_find_ctx_download_merged: no sharded launch on this context");
    if (set_device(c->st)) return -1;
    const int k = c->merged_topk, count = c->merged_count;
    MergedLayout L(k, count);
    if (c->h_out.reserve(L.total)) return -1;
    uint8_t* h = (uint8_t*)c->h_out.p;
    CU_OK(cudaMemcpyAsync(h, c->merged.p, L.total, cudaMemcpyDeviceToHost, c->stream));
    CU_OK(cudaStreamSynchronize(c->stream));
    for (int q = 0; q < count; q++) {
        const int m = ((const int*)(h + L.n))[q];
        const size_t o = (size_t)q * k;
        if (out_n) out_n[q] = m;
        if (out_ids) memcpy(out_ids + o, h + L.ids + o * 8, (size_t)m * 8);
        if (out_sims) memcpy(out_sims + o, h + L.sims + o * 4, (size_t)m * 4);
        if (out_shard) memcpy(out_shard + o, h + L.shard + o * 4, (size_t)m * 4);
        if (out_index) memcpy(out_index + o, h + L.index + o * 4, (size_t)m * 4);
    }
    return 0;
}

extern "C" int tfp_find_ctx_download_merged_keys(tfp_find_ctx* c, int* out_n, int64_t* out_ids, float* out_sims,
                                                 float* out_keys) {
    if (!c || c->merged_topk <= 0 || c->merged_keys_dim <= 0)
        return fail(2, "tfp_find_ctx_download_merged_keys: no sharded launch with keys on this context");
    if (set_device(c->st)) return -1;
    const int k = c->merged_topk, count = c->merged_count, dim = c->merged_keys_dim;
    MergedLayout L(k, count);
    const size_t total = L.total + (size_t)count * k * dim * 4;
    if (c->h_out.reserve(total)) return -1;
    uint8_t* h = (uint8_t*)c->h_out.p;
    CU_OK(cudaMemcpyAsync(h, c->merged.p, total, cudaMemcpyDeviceToHost, c->stream));   // ONE copy: n, ids, sims, rows
    CU_OK(cudaStreamSynchronize(c->stream));
    for (int q = 0; q < count; q++) {
        const int m = ((const int*)(h + L.n))[q];
        const size_t o = (size_t)q * k;
        if (out_n) out_n[q] = m;
        if (out_ids) memcpy(out_ids + o, h + L.ids + o * 8, (size_t)m * 8);
        if (out_sims) memcpy(out_sims + o, h + L.sims + o * 4, (size_t)m * 4);
        if (out_keys) memcpy(out_keys + o * dim, h + L.total + o * dim * 4, (size_t)m * dim * 4);
    }
    return 0;
}

// ---------------------------------------------------------------------------
// pinned host staging
// ---------------------------------------------------------------------------
extern "C" void* tfp_host_alloc(size_t bytes) {
    void* p = nullptr;
    cudaError_t e = cudaMallocHost(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) { fail(-1, "cudaMallocHost(%zu) failed: %s", bytes, cudaGetErrorString(e)); return nullptr; }
    return p;
}
extern "C" void tfp_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---------------------------------------------------------------------------
// one logical store over several GPUs in one process
// ---------------------------------------------------------------------------
#include "multi_store.inl"

// ---------------------------------------------------------------------------
// The store's public entry points: dispatch on the kind of handle, keep the error text per handle.
// ---------------------------------------------------------------------------
static int keep_err(tfp_store* s, int rc) {
    if (rc && s) {
        std::lock_guard<std::mutex> g(s->err_mu);
        s->last_err = g_err;
    }
    return rc;
}

extern "C" int tfp_store_open(const tfp_cfg* cfg, tfp_store** out) { return store_open(cfg, out); }

extern "C" int tfp_store_open_devices(const tfp_cfg* cfg, const int* devices, int n_devices, tfp_store** out) {
    if (!out) return fail(2, "tfp_store_open_devices: out is NULL");
    *out = nullptr;
    if (!devices || n_devices < 1) return fail(2, "tfp_store_open_devices: no devices");
    if (n_devices == 1) {
        tfp_cfg c;
        memset(&c, 0, sizeof(c));
        if (cfg) c = *cfg;
        c.device = devices[0];
        return store_open(&c, out);
    }
    return multi_open(cfg, devices, n_devices, out);
}

extern "C" void tfp_store_close(tfp_store* s) {
    if (!s) return;
    if (s->multi) { multi_close(s); delete s; return; }
    store_close(s);
}

extern "C" int tfp_store_last_error(tfp_store* s, char* buf, size_t cap) {
    if (!s || !buf || cap == 0) return fail(2, "tfp_store_last_error: bad argument");
    std::lock_guard<std::mutex> g(s->err_mu);
    snprintf(buf, cap, "%s", s->last_err.c_str());
    return 0;
}

extern "C" int tfp_store_shard_count(tfp_store* s) { return !s ? 0 : (s->multi ? (int)s->multi->shards.size() : 1); }

extern "C" int tfp_store_shard_stats(tfp_store* s, int shard, tfp_stats* out) {
    if (!s || !out) return fail(2, "tfp_store_shard_stats: NULL argument");
    if (shard < 0 || shard >= tfp_store_shard_count(s)) return keep_err(s, fail(2, "tfp_store_shard_stats: shard %d out of range", shard));
    return keep_err(s, store_stats(s->multi ? s->multi->shards[(size_t)shard] : s, out));
}

extern "C" int tfp_route_rows(const float* keys, int64_t n, int dim, int n_shards, int* out_shard) {
    if (!keys || !out_shard || n < 0 || dim <= 0 || n_shards < 1) return fail(2, "tfp_route_rows: bad argument");
    for (int64_t i = 0; i < n; i++) out_shard[i] = route_of(keys + (size_t)i * dim, dim, n_shards);
    return 0;
}

extern "C" int tfp_store_set(tfp_store* s, const float* keys, int64_t n, int dim, int64_t n_values, int64_t* out_ids,
                             uint8_t* out_was_new) {
    if (!s) return fail(2, "tfp_store_set: store is NULL");
    return keep_err(s, s->multi ? multi_set(s, keys, n, dim, n_values, out_ids, out_was_new)
                                : store_set(s, keys, n, dim, n_values, out_ids, out_was_new));
}
extern "C" int tfp_store_set_device(tfp_store* s, const float* d_keys, int64_t n, int dim, int64_t* out_ids,
                                    uint8_t* out_was_new) {
    if (!s) return fail(2, "tfp_store_set_device: store is NULL");
    if (s->multi) return keep_err(s, multi_set_device(s, d_keys, n, dim, out_ids, out_was_new));
    return keep_err(s, store_set_device(s, d_keys, n, dim, out_ids, out_was_new));
}
extern "C" int tfp_store_delete(tfp_store* s, const float* keys, int64_t n, int dim, int64_t* out_ids) {
    if (!s) return fail(2, "tfp_store_delete: store is NULL");
    return keep_err(s, s->multi ? multi_delete(s, keys, n, dim, out_ids) : store_delete(s, keys, n, dim, out_ids));
}
extern "C" int tfp_store_get(tfp_store* s, const float* keys, int64_t n, int dim, int64_t* out_ids, int64_t* out_order,
                             int64_t* out_n_found) {
    if (!s) return fail(2, "tfp_store_get: store is NULL");
    return keep_err(s, s->multi ? multi_get(s, keys, n, dim, out_ids, out_order, out_n_found)
                                : store_get(s, keys, n, dim, out_ids, out_order, out_n_found));
}
extern "C" int tfp_store_find_batch(tfp_store* s, const float* q, int nq, int dim, int topk, int* out_n, int64_t* out_ids,
                                    float* out_sims, float* out_keys) {
    if (!s) return fail(2, "tfp_store_find: store is NULL");
    return keep_err(s, s->multi ? multi_find_batch(s, q, nq, dim, topk, out_n, out_ids, out_sims, out_keys, nullptr)
                                : store_find_batch(s, q, nq, dim, topk, out_n, out_ids, out_sims, out_keys, nullptr));
}
extern "C" int tfp_store_find(tfp_store* s, const float* q, int dim, int topk, int* out_n, int64_t* out_ids,
                              float* out_sims, float* out_keys, int* out_path) {
    if (!s) return fail(2, "tfp_store_find: store is NULL");
    return keep_err(s, s->multi ? multi_find_batch(s, q, 1, dim, topk, out_n, out_ids, out_sims, out_keys, out_path)
                                : store_find_batch(s, q, 1, dim, topk, out_n, out_ids, out_sims, out_keys, out_path));
}
extern "C" int tfp_store_stats(tfp_store* s, tfp_stats* out) {
    if (!s || !out) return fail(2, "tfp_store_stats: NULL argument");
    return keep_err(s, s->multi ? multi_stats(s, out) : store_stats(s, out));
}
extern "C" int tfp_store_mark_not_normalized(tfp_store* s) {
    if (!s) return fail(2, "tfp_store_mark_not_normalized: store is NULL");
    return keep_err(s, s->multi ? multi_mark_not_normalized(s) : store_mark_not_normalized(s));
}
extern "C" int tfp_store_save_tagged(tfp_store* s, const char* path, uint64_t tag) {
    if (!s || !path) return fail(2, "tfp_store_save: NULL argument");
    return keep_err(s, s->multi ? multi_save(s, path, tag) : store_save(s, path, tag));
}
extern "C" int tfp_store_save(tfp_store* s, const char* path) { return tfp_store_save_tagged(s, path, 0); }
extern "C" int tfp_snapshot_tag(const char* path, uint64_t* out_tag) {
    if (!path || !out_tag) return fail(2, "tfp_snapshot_tag: NULL argument");
    *out_tag = 0;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(2, "tfp_snapshot_tag: cannot open %s", path);
    char magic[8];
    bool ok = fread(magic, 8, 1, f) == 1;
    if (ok && memcmp(magic, "TFPSTOR2", 8) == 0) ok = fseek(f, (long)kSnapHeaderV1Bytes, SEEK_SET) == 0 && fread(out_tag, 8, 1, f) == 1;
    else if (ok && memcmp(magic, "TFPMULT1", 8) == 0) ok = fseek(f, (long)offsetof(MultiManifest, tag), SEEK_SET) == 0 && fread(out_tag, 8, 1, f) == 1;
    else if (ok && memcmp(magic, "TFPSTOR1", 8) == 0) *out_tag = 0;
    else ok = false;
    fclose(f);
    return ok ? 0 : fail(2, "tfp_snapshot_tag: %s is not a store snapshot", path);
}
extern "C" int tfp_store_load(tfp_store* s, const char* path) {
    if (!s || !path) return fail(2, "tfp_store_load: NULL argument");
    return keep_err(s, s->multi ? multi_load(s, path) : store_load(s, path));
}
