// multi_store.inl -- ONE logical store over several GPUs inside ONE process.
//
// LocalAI spawns exactly one `local-store` process per backend and hands it nothing but `--addr` and the
// environment (/root/reference/pkg/model/initializers.go:333-368, pkg/model/process.go:94-124), so the
// multi-GPU store has to live behind the same handle and the same entry points as the one-GPU store:
// tfp_store_open_devices returns a tfp_store* that tfp_store_set / _delete / _get / _find / _find_batch /
// _stats / _save / _load / _close serve like any other.
//
//   * G shards, each a complete one-GPU store (its own HBM arrays, host index, find contexts).
//   * Keys are routed to their owner shard by a hash of a few canonical key elements (route_of), so the
//     exact-match operations of store.go:106-311 touch one shard per key.
//   * One worker thread per shard does that shard's CUDA calls (cudaSetDevice is per OS thread; a Go host's
//     goroutines migrate between threads); shard 0's work runs on the calling thread.
//   * A find: every shard scans its rows; the scan kernel's last CTA writes the shard's fixed-size result
//     record straight into shard 0's gather block over NVLink (plain peer pointers: cudaDeviceEnablePeerAccess)
//     and raises a sequence word there; ONE merge kernel on GPU 0 waits for the words and merges; one D2H copy
//     returns the answer.  No collective library, no cudaIpc, no torch.  Several host threads can be inside
//     tfp_store_find at once: each takes one of the store's lanes (a set of per-shard find contexts plus a
//     gather block), so finds overlap on the GPUs.
//   * Answers too large for the gather block (huge top-k, or key rows of a big batch) take the general path:
//     per-shard finds and a k-way merge on the host.
//
// Included at the end of tfp_store.cu (it uses that file's internals).

#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <thread>
#include <immintrin.h>

namespace {

constexpr int kMultiMaxShards = kXchgMaxWorld;
constexpr int kMultiIdShift = 40;          // shard g hands out ids from id_base + (g << 40), like the per-process sharding

// Owner shard of a key: a hash of up to 16 elements spread over the row, in canonical bits, so that keys
// equal under Go's cmp.Compare (-0 == +0, NaN == NaN; store.go:48-52) always meet in the same shard.
// O(1) per key: the host never walks whole rows (1M keys/s x 6 KiB would not fit one core).
inline int route_of(const float* row, int dim, int world) {
    uint64_t h = 0x9E3779B97F4A7C15ull ^ (uint64_t)(uint32_t)dim;
    const int samples = dim < 16 ? dim : 16;
    for (int j = 0; j < samples; j++) {
        const int i = (int)((int64_t)j * dim / samples);
        h = mix64(h ^ (((uint64_t)(uint32_t)i << 32) | (uint64_t)canon_bits(row[i])));
    }
    return (int)(h % (uint64_t)world);
}

struct Job {
    std::function<int()> fn;
    int rc = 0;
    std::string err;
    std::atomic<int> done{0};
};

struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<Job*> q;
    std::atomic<int> pending{0};
    bool stop = false;

    void loop() {
        for (;;) {
            // a find is tens of microseconds: poll briefly before sleeping, a condition-variable wake-up costs about as much
            for (int spin = 0; spin < 20000 && pending.load(std::memory_order_acquire) == 0; spin++) _mm_pause();
            Job* j = nullptr;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || !q.empty(); });
                if (q.empty()) return;
                j = q.front();
                q.pop_front();
                pending.fetch_sub(1, std::memory_order_relaxed);
            }
            j->rc = j->fn();
            if (j->rc) j->err = g_err;
            j->done.store(1, std::memory_order_release);
        }
    }
    void post(Job* j) {
        {
            std::lock_guard<std::mutex> lk(mu);
            q.push_back(j);
            pending.fetch_add(1, std::memory_order_release);
        }
        cv.notify_one();
    }
    void shutdown() {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
        }
        cv.notify_one();
        if (th.joinable()) th.join();
    }
};

inline void job_wait(Job* j) {
    for (int spin = 0; j->done.load(std::memory_order_acquire) == 0; spin++) {
        if (spin < 50000) _mm_pause();
        else std::this_thread::sleep_for(std::chrono::microseconds(20));
    }
}

struct MultiLane {
    std::vector<tfp_find_ctx*> ctx;   // one find context per shard; ctx[0] owns the gather block
    bool busy = false;
};

}  // namespace

struct tfp_multi {
    std::vector<tfp_store*> shards;
    std::vector<int> devices;
    std::vector<Worker*> workers;     // workers[g] serves shard g (g >= 1)
    std::vector<MultiLane> lanes;
    std::mutex lane_mu;
    std::condition_variable lane_cv;
    size_t slot_bytes = 0;
    bool p2p = false;                 // every shard can write into shard 0's memory (same GPU, or peer access)
    bool shared_device = false;       // two shards sit on one GPU (a test configuration)
    std::vector<PinBuf> stage;        // per shard: page-locked staging of the rows routed to it
};

// Run fn(g) for every shard in `which`: the first on the calling thread, the others on their workers, all at once.
// Returns the first failure (its text becomes the calling thread's error text).
template <class F>
static int run_shards(tfp_multi* m, const std::vector<int>& which, F&& fn) {
    if (which.empty()) return 0;
    // `which` is ascending, so only which[0] can be shard 0 (which has no worker: it runs here, like any first shard)
    std::vector<Job> jobs(which.size());
    for (size_t i = 1; i < which.size(); i++) {
        const int g = which[i];
        jobs[i].fn = [&fn, g]() -> int { return fn(g); };
        m->workers[(size_t)g]->post(&jobs[i]);
    }
    int rc = fn(which[0]);
    std::string err = rc ? g_err : std::string();
    for (size_t i = 1; i < which.size(); i++) {
        job_wait(&jobs[i]);
        if (jobs[i].rc && !rc) { rc = jobs[i].rc; err = jobs[i].err; }
    }
    if (rc) g_err = err;
    return rc;
}

static std::vector<int> all_shards(const tfp_multi* m) {
    std::vector<int> v(m->shards.size());
    for (size_t g = 0; g < v.size(); g++) v[g] = (int)g;
    return v;
}

static void multi_close(tfp_store* top) {
    tfp_multi* m = top->multi;
    if (!m) return;
    if (tuning().set_trace && top->tr_calls > 0)
        fprintf(stderr, "tfp multi set trace (%lld calls, us per call): wait for the mutation lock %.0f | routing %.0f | slowest shard's gather %.0f | "
                        "all shards %.0f\n", (long long)top->tr_calls, top->tr_us[3] / top->tr_calls, top->tr_us[0] / top->tr_calls,
                top->tr_us[1] / top->tr_calls, top->tr_us[2] / top->tr_calls);
    if (tuning().set_trace && top->trd_calls > 0)
        fprintf(stderr, "tfp multi delete trace (%lld calls, us per call): wait for the mutation lock %.0f | the call %.0f\n",
                (long long)top->trd_calls, top->trd_us[0] / top->trd_calls, top->trd_us[1] / top->trd_calls);
    for (Worker* w : m->workers) if (w) { w->shutdown(); delete w; }
    for (size_t g = 0; g < m->shards.size(); g++) {
        if (!m->shards[g]) continue;
        cudaSetDevice(m->shards[g]->device);
        if (g < m->stage.size()) m->stage[g].release();
        store_close(m->shards[g]);   // destroys the shard's find contexts (the lanes') too
    }
    delete m;
    top->multi = nullptr;
}

static int multi_open(const tfp_cfg* cfg, const int* devices, int n_devices, tfp_store** out) {
    if (n_devices < 2 || n_devices > kMultiMaxShards)
        return fail(2, "tfp_store_open_devices: %d devices (one logical store spans 1..%d GPUs)", n_devices, kMultiMaxShards);
    tfp_store* top = new tfp_store();
    tfp_multi* m = new tfp_multi();
    top->multi = m;
    top->device = devices[0];
    m->devices.assign(devices, devices + n_devices);
    m->shards.assign((size_t)n_devices, nullptr);
    m->workers.assign((size_t)n_devices, nullptr);
    m->stage.resize((size_t)n_devices);
    auto bail = [&](int rc) { std::string e = g_err; multi_close(top); delete top; g_err = e; return rc; };
    for (int g = 0; g < n_devices; g++) {
        tfp_cfg c;
        memset(&c, 0, sizeof(c));
        c.device = devices[g];
        c.capacity_rows = cfg ? (cfg->capacity_rows + n_devices - 1) / n_devices : 0;
        c.flags = cfg ? cfg->flags : 0;
        c.id_base = (cfg ? cfg->id_base : 0) + ((int64_t)g << kMultiIdShift);
        int rc = store_open(&c, &m->shards[(size_t)g]);
        if (rc) return bail(rc);
    }
    for (int g = 0; g < n_devices; g++)
        for (int h = 0; h < g; h++) if (devices[g] == devices[h]) m->shared_device = true;
    // every shard must be able to write into shard 0's gather blocks
    m->p2p = true;
    for (int g = 1; g < n_devices; g++) {
        if (devices[g] == devices[0]) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, devices[g], devices[0]) != cudaSuccess || !can) { m->p2p = false; cudaGetLastError(); continue; }
        if (cudaSetDevice(devices[g]) != cudaSuccess) { m->p2p = false; cudaGetLastError(); continue; }
        cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) m->p2p = false;
        cudaGetLastError();
    }
    size_t slot = (size_t)1 << 20;
    if (const char* e = getenv("TFP_MULTI_SLOT_BYTES")) slot = (size_t)atoll(e);
    slot = std::min<size_t>(std::max<size_t>(slot, 4096), (size_t)64 << 20);
    slot = (slot + 255) & ~(size_t)255;
    m->slot_bytes = slot;
    int n_lanes = 4;
    if (const char* e = getenv("TFP_MULTI_LANES")) n_lanes = std::min(16, std::max(1, atoi(e)));
    m->lanes.resize((size_t)n_lanes);
    for (MultiLane& L : m->lanes) {
        L.ctx.assign((size_t)n_devices, nullptr);
        for (int g = 0; g < n_devices; g++) {
            tfp_store* sh = m->shards[(size_t)g];
            if (set_device(sh)) return bail(-1);
            if (ctx_create(sh, &L.ctx[(size_t)g])) return bail(-1);
        }
        tfp_find_ctx* root = L.ctx[0];
        if (set_device(m->shards[0])) return bail(-1);
        const size_t total = xbuf_flags_offset(n_devices, slot) + 256;
        if (root->xbuf.reserve(total)) return bail(-1);
        if (cudaMemset(root->xbuf.p, 0, total) != cudaSuccess) return bail(fail(-1, "cudaMemset failed"));
        for (int g = 0; g < n_devices; g++) {
            tfp_find_ctx* c = L.ctx[(size_t)g];
            c->x_root_mode = true;
            c->x_world = n_devices;
            c->x_me = g;
            c->x_slot = slot;
            c->x_seq = 0;
            c->x_peer[0] = root->xbuf.as<uint8_t>();
            c->x_ready = m->p2p;
            if (m->shared_device) {
                // Two shards on one GPU (a test configuration): a merge kernel that spins until the other shard's scan
                // has run would be one launch waiting for another on the SAME GPU, and nothing guarantees that the two
                // run at the same time.  There the merge is ordered behind the scans with events instead.
                c->x_no_spin = true;
                if (set_device(m->shards[(size_t)g])) return bail(-1);
                if (cudaEventCreateWithFlags(&c->x_done_ev, cudaEventDisableTiming) != cudaSuccess) return bail(fail(-1, "cudaEventCreate failed"));
            }
        }
    }
    for (int g = 1; g < n_devices; g++) {
        Worker* w = new Worker();
        m->workers[(size_t)g] = w;
        w->th = std::thread([w] { w->loop(); });
    }
    *out = top;
    return 0;
}

// store.go:115-116 / :209-210 / :275-276: the first Set, Delete or Get fixes the key length -- of every shard, so
// that a shard which owns no key yet answers finds with an empty list instead of a length error.
static int multi_latch(tfp_store* top, int dim, const char* fmt) {
    {   // the usual case -- the length is fixed and matches -- needs no exclusive lock (finds hold this lock shared
        // for their whole duration: an exclusive request per StoresSet would queue behind all of them)
        std::shared_lock<RwMutex> lk(top->rw);
        if (top->dim == dim) return 0;
    }
    std::unique_lock<RwMutex> lk(top->rw);
    if (top->dim == -1) {
        top->dim = dim;
        for (tfp_store* sh : top->multi->shards) {
            std::unique_lock<RwMutex> lk(sh->rw);
            if (sh->dim == -1) sh->dim = dim;
        }
    } else if (dim != top->dim) {
        return fail(1, fmt, dim, top->dim);
    }
    return 0;
}

struct Routed {
    std::vector<std::vector<int64_t>> idx;   // per shard: request indices, in request order
    std::vector<int> active;                 // shards with at least one row, shard 0 first if present
};
static void route_rows(const tfp_multi* m, const float* keys, int64_t n, int dim, Routed& r) {
    const int G = (int)m->shards.size();
    r.idx.assign((size_t)G, {});
    // The 16 sampled elements of a row sit in 16 different cache lines of memory nobody has touched yet: without
    // help the loop runs at one DRAM latency per sample (measured 2.1 ms for 8192 rows of 1536 floats, as long as
    // everything the GPUs do for the call).  Prefetch the samples of the row a few iterations ahead.
    const int samples = dim < 16 ? dim : 16;
    constexpr int kAhead = 6;
    for (int64_t i = 0; i < n; i++) {
        if (i + kAhead < n) {
            const float* nxt = keys + (size_t)(i + kAhead) * dim;
            for (int j = 0; j < samples; j++) __builtin_prefetch(nxt + (int64_t)j * dim / samples, 0, 0);
        }
        r.idx[(size_t)route_of(keys + (size_t)i * dim, dim, G)].push_back(i);
    }
    for (int g = 0; g < G; g++) if (!r.idx[(size_t)g].empty()) r.active.push_back(g);
}

// Copy the rows routed to shard g into its page-locked staging buffer (on the thread that serves the shard).
static int stage_routed(tfp_multi* m, int g, const float* keys, int dim, const std::vector<int64_t>& idx, const float** out) {
    if (set_device(m->shards[(size_t)g])) return -1;
    const size_t row = (size_t)dim * 4;
    if (m->stage[(size_t)g].reserve(idx.size() * row)) return -1;
    uint8_t* dst = m->stage[(size_t)g].as<uint8_t>();
    for (size_t j = 0; j < idx.size(); j++) memcpy(dst + j * row, keys + (size_t)idx[j] * dim, row);
    *out = (const float*)dst;
    return 0;
}

static int multi_set(tfp_store* top, const float* keys, int64_t n, int dim, int64_t n_values, int64_t* out_ids,
                     uint8_t* out_was_new) {
    tfp_multi* m = top->multi;
    if (n == 0) return fail(1, "no keys to add");                                     // store.go:107-109
    if (n != n_values) return fail(1, "len(keys) = %lld, len(values) = %lld", (long long)n, (long long)n_values);
    if (!keys || dim <= 0) return fail(2, "tfp_store_set: keys is NULL or dim <= 0");
    if (check_scatter_grid(n)) return 2;
    const double t_wait = now_us();
    std::lock_guard<std::mutex> mut(top->mut_mu);   // one mutation at a time; finds go on (the shards publish appends themselves)
    if (multi_latch(top, dim, "Try to add key with length %d when existing length is %d")) return 1;
    const double t0 = now_us();
    top->tr_us[3] += t0 - t_wait;
    Routed r;
    route_rows(m, keys, n, dim, r);
    const double t1 = now_us();
    std::vector<std::vector<int64_t>> ids(m->shards.size());
    std::vector<std::vector<uint8_t>> isnew(m->shards.size());
    std::vector<double> t_stage(m->shards.size(), 0.0);
    int rc = run_shards(m, r.active, [&](int g) -> int {
        const std::vector<int64_t>& idx = r.idx[(size_t)g];
        const float* rows = nullptr;
        const double a = now_us();
        if (stage_routed(m, g, keys, dim, idx, &rows)) return -1;
        t_stage[(size_t)g] = now_us() - a;
        ids[(size_t)g].resize(idx.size());
        isnew[(size_t)g].resize(idx.size());
        return store_set(m->shards[(size_t)g], rows, (int64_t)idx.size(), dim, (int64_t)idx.size(), ids[(size_t)g].data(),
                         isnew[(size_t)g].data());
    });
    if (rc) return rc;
    top->tr_us[0] += t1 - t0;                                             // routing
    top->tr_us[1] += *std::max_element(t_stage.begin(), t_stage.end());   // slowest shard's gather into pinned staging
    top->tr_us[2] += now_us() - t1;                                       // all shards, start to finish
    top->tr_calls++;
    for (int g : r.active)
        for (size_t j = 0; j < r.idx[(size_t)g].size(); j++) {
            const int64_t i = r.idx[(size_t)g][j];
            if (out_ids) out_ids[i] = ids[(size_t)g][j];
            if (out_was_new) out_was_new[i] = isnew[(size_t)g][j];
        }
    // the sticky flag (store.go:27,126-127) belongs to the whole store: all shards leave the normalised path together
    bool all_norm = true;
    for (tfp_store* sh : m->shards) { std::shared_lock<RwMutex> l2(sh->rw); all_norm = all_norm && sh->normalized; }
    if (!all_norm && top->normalized) {   // (only the call that flips the flag pays for the exclusive lock: it waits for every find in flight)
        std::unique_lock<RwMutex> lk(top->rw);
        top->normalized = false;
        for (tfp_store* sh : m->shards) store_mark_not_normalized(sh);
    }
    return 0;
}

// Keys that already sit in one GPU's memory.  They still have to be routed, and routing looks at the rows: the block is
// brought to page-locked host memory in chunks and goes through multi_set -- correct and simple, one PCIe round trip
// (a device-side route + peer gather would save it; not built).
static int multi_set_device(tfp_store* top, const float* d_keys, int64_t n, int dim, int64_t* out_ids, uint8_t* out_was_new) {
    if (n == 0) return fail(1, "no keys to add");
    if (!d_keys || dim <= 0) return fail(2, "tfp_store_set_device: d_keys is NULL or dim <= 0");
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, d_keys) != cudaSuccess || at.type != cudaMemoryTypeDevice) {
        cudaGetLastError();
        return fail(2, "tfp_store_set_device: d_keys is not device memory");
    }
    CU_OK(cudaSetDevice(at.device));
    CU_OK(cudaDeviceSynchronize());   // the caller's producer stream is unknown: order after all of it
    const int64_t chunk_rows = std::max<int64_t>(1, std::min<int64_t>(n, (int64_t)(((size_t)256 << 20) / ((size_t)dim * 4))));
    PinBuf pin;
    if (pin.reserve((size_t)chunk_rows * dim * 4)) return -1;
    int rc = 0;
    for (int64_t lo = 0; lo < n && !rc; lo += chunk_rows) {
        const int64_t m = std::min(chunk_rows, n - lo);
        if (cudaSetDevice(at.device) != cudaSuccess ||
            cudaMemcpy(pin.p, d_keys + (size_t)lo * dim, (size_t)m * dim * 4, cudaMemcpyDeviceToHost) != cudaSuccess) {
            rc = fail(-1, "tfp_store_set_device: device-to-host copy failed");
            break;
        }
        rc = multi_set(top, pin.as<float>(), m, dim, m, out_ids ? out_ids + lo : nullptr, out_was_new ? out_was_new + lo : nullptr);
    }
    const std::string err = g_err;
    pin.release();
    if (rc) g_err = err;
    return rc;
}

static int multi_delete(tfp_store* top, const float* keys, int64_t n, int dim, int64_t* out_ids) {
    tfp_multi* m = top->multi;
    if (n == 0) return fail(1, "no keys to delete");                                 // store.go:201-203
    if (!keys || dim <= 0) return fail(2, "tfp_store_delete: keys is NULL or dim <= 0");
    const double td_wait = now_us();
    std::lock_guard<std::mutex> mut(top->mut_mu);
    const double td0 = now_us();
    if (multi_latch(top, dim, "Trying to delete key with length %d when existing length is %d")) return 1;
    if (out_ids) for (int64_t i = 0; i < n; i++) out_ids[i] = -1;
    Routed r;
    route_rows(m, keys, n, dim, r);
    std::vector<std::vector<int64_t>> ids(m->shards.size());
    int rc = run_shards(m, r.active, [&](int g) -> int {
        const std::vector<int64_t>& idx = r.idx[(size_t)g];
        const float* rows = nullptr;
        if (stage_routed(m, g, keys, dim, idx, &rows)) return -1;
        ids[(size_t)g].assign(idx.size(), -1);
        return store_delete(m->shards[(size_t)g], rows, (int64_t)idx.size(), dim, ids[(size_t)g].data());
    });
    if (rc) return rc;
    if (out_ids)
        for (int g : r.active)
            for (size_t j = 0; j < r.idx[(size_t)g].size(); j++) out_ids[r.idx[(size_t)g][j]] = ids[(size_t)g][j];
    top->trd_us[0] += td0 - td_wait;      // wait for the mutation lock
    top->trd_us[1] += now_us() - td0;     // the call itself
    top->trd_calls++;
    return 0;
}

static int multi_get(tfp_store* top, const float* keys, int64_t n, int dim, int64_t* out_ids, int64_t* out_order,
                     int64_t* out_n_found) {
    tfp_multi* m = top->multi;
    if (out_n_found) *out_n_found = 0;
    if (n == 0) return fail(1, "no keys to get");   // reference panics here (store.go:276); documented deviation
    if (!keys || dim <= 0 || !out_ids) return fail(2, "tfp_store_get: NULL argument or dim <= 0");
    std::lock_guard<std::mutex> mut(top->mut_mu);
    if (multi_latch(top, dim, "Try to get a key with length %d when existing length is %d")) return 1;
    for (int64_t i = 0; i < n; i++) out_ids[i] = -1;
    Routed r;
    route_rows(m, keys, n, dim, r);
    std::vector<std::vector<int64_t>> ids(m->shards.size());
    int rc = run_shards(m, r.active, [&](int g) -> int {
        const std::vector<int64_t>& idx = r.idx[(size_t)g];
        const float* rows = nullptr;
        if (stage_routed(m, g, keys, dim, idx, &rows)) return -1;
        ids[(size_t)g].assign(idx.size(), -1);
        int64_t found = 0;
        return store_get(m->shards[(size_t)g], rows, (int64_t)idx.size(), dim, ids[(size_t)g].data(), nullptr, &found);
    });
    if (rc) return rc;
    for (int g : r.active)
        for (size_t j = 0; j < r.idx[(size_t)g].size(); j++) out_ids[r.idx[(size_t)g][j]] = ids[(size_t)g][j];
    // found keys in the reference's sorted-key order (store.go:269,285-301); equal keys keep request order
    std::vector<int64_t> hit;
    for (int64_t i = 0; i < n; i++) if (out_ids[i] >= 0) hit.push_back(i);
    if (out_order) {
        std::stable_sort(hit.begin(), hit.end(), [&](int64_t a, int64_t b) {
            return cmp_rows(keys + (size_t)a * dim, keys + (size_t)b * dim, dim) < 0;
        });
        for (size_t j = 0; j < hit.size(); j++) out_order[j] = hit[j];
    }
    if (out_n_found) *out_n_found = (int64_t)hit.size();
    return 0;
}

static int multi_stats(tfp_store* top, tfp_stats* out) {
    tfp_multi* m = top->multi;
    std::shared_lock<RwMutex> lk(top->rw);
    memset(out, 0, sizeof(*out));
    for (tfp_store* sh : m->shards) {
        tfp_stats st;
        int rc = store_stats(sh, &st);
        if (rc) return rc;
        out->rows += st.rows; out->capacity_rows += st.capacity_rows; out->hbm_bytes += st.hbm_bytes;
    }
    out->dim = top->dim;
    out->keys_are_normalized = top->normalized ? 1 : 0;
    out->device = m->devices[0];
    return 0;
}

static int multi_mark_not_normalized(tfp_store* top) {
    std::unique_lock<RwMutex> lk(top->rw);
    top->normalized = false;
    for (tfp_store* sh : top->multi->shards) store_mark_not_normalized(sh);
    return 0;
}

static int multi_lane_acquire(tfp_multi* m) {
    std::unique_lock<std::mutex> lk(m->lane_mu);
    for (;;) {
        for (size_t i = 0; i < m->lanes.size(); i++)
            if (!m->lanes[i].busy) { m->lanes[i].busy = true; return (int)i; }
        m->lane_cv.wait(lk);
    }
}
static void multi_lane_release(tfp_multi* m, int lane) {
    { std::lock_guard<std::mutex> lk(m->lane_mu); m->lanes[(size_t)lane].busy = false; }
    m->lane_cv.notify_one();
}

// The device path: scans on every GPU, records (and, when asked for, the winners' key rows) pushed into shard 0's
// gather block by the scan kernels, one waiting merge kernel on GPU 0, one D2H copy.
static int multi_find_device(tfp_multi* m, MultiLane& L, const float* q, int nq, int dim, int topk, const QInfo* qi,
                             int* out_n, int64_t* out_ids, float* out_sims, float* out_keys, int* out_path) {
    const bool with_keys = out_keys != nullptr;
    const int G = (int)m->shards.size();
    std::vector<int> launched((size_t)G, 0);
    int rc = run_shards(m, all_shards(m), [&](int g) -> int {
        tfp_store* sh = m->shards[(size_t)g];
        tfp_find_ctx* c = L.ctx[(size_t)g];
        std::shared_lock<RwMutex> lk(sh->rw);
        if (set_device(sh)) return -1;
        int r = ctx_upload_locked(c, q, nq, dim, qi);
        if (r) return r;
        r = p2p_launch_locked(c, 0, nq, topk, with_keys, /*defer_merge=*/true);
        if (r) return r;
        if (c->x_done_ev) CU_OK(cudaEventRecord(c->x_done_ev, c->stream));
        launched[(size_t)g] = 1;
        return 0;
    });
    tfp_find_ctx* root = L.ctx[0];
    if (rc) {
        // A shard could not launch: no merge is enqueued (nothing waits for the missing records); bring the lane's
        // sequence numbers back together so that the next find on it pairs up again.
        unsigned int top_seq = 0;
        for (int g = 0; g < G; g++) top_seq = std::max(top_seq, L.ctx[(size_t)g]->x_seq);
        for (int g = 0; g < G; g++) L.ctx[(size_t)g]->x_seq = top_seq;
        return rc;
    }
    // every shard's scan is in its stream: now the merge kernel, which waits for their records, can go in
    if (set_device(m->shards[0])) return -1;
    if (root->x_no_spin)
        for (int g = 1; g < G; g++) CU_OK(cudaStreamWaitEvent(root->stream, L.ctx[(size_t)g]->x_done_ev, 0));
    {
        std::shared_lock<RwMutex> lk(m->shards[0]->rw);
        rc = p2p_merge_locked(root);
        if (rc) return rc;
    }
    if (out_path) *out_path = root->last_path;
    return with_keys ? tfp_find_ctx_download_merged_keys(root, out_n, out_ids, out_sims, out_keys)
                     : tfp_find_ctx_download_merged(root, out_n, out_ids, out_sims, nullptr, nullptr);
}

// The general path: a complete find on every shard, merged on the host (similarity descending; among equal
// similarities the lower shard, then the shard's own order -- what the merge kernel does).
static int multi_find_general(tfp_multi* m, MultiLane& L, const float* q, int nq, int dim, int topk, const QInfo* qi,
                              int* out_n, int64_t* out_ids, float* out_sims, float* out_keys, int* out_path) {
    const int G = (int)m->shards.size();
    struct Part { int k = 0; std::vector<int> n; std::vector<int64_t> ids; std::vector<float> sims, keys; };
    std::vector<Part> parts((size_t)G);
    int rc = run_shards(m, all_shards(m), [&](int g) -> int {
        tfp_store* sh = m->shards[(size_t)g];
        tfp_find_ctx* c = L.ctx[(size_t)g];
        std::shared_lock<RwMutex> lk(sh->rw);
        if (set_device(sh)) return -1;
        Part& P = parts[(size_t)g];
        P.k = (int)std::min<int64_t>((int64_t)topk, std::max<int64_t>(sh->n, 1));   // a shard cannot return more than it holds
        P.n.assign((size_t)nq, 0);
        P.ids.resize((size_t)nq * P.k);
        P.sims.resize((size_t)nq * P.k);
        if (out_keys) P.keys.resize((size_t)nq * P.k * dim);
        int r = ctx_upload_locked(c, q, nq, dim, qi);
        if (r) return r;
        r = ctx_launch_locked(c, 0, nq, P.k, nullptr, out_keys != nullptr, /*local_records=*/true);
        if (r) return r;
        return ctx_download_locked(c, P.n.data(), P.ids.data(), P.sims.data(), out_keys ? P.keys.data() : nullptr);
    });
    if (rc) return rc;
    if (out_path) *out_path = L.ctx[0]->last_path;
    std::vector<int> pos((size_t)G);
    for (int qq = 0; qq < nq; qq++) {
        std::fill(pos.begin(), pos.end(), 0);
        int filled = 0;
        while (filled < topk) {
            int best = -1;
            uint32_t best_o = 0;
            for (int g = 0; g < G; g++) {
                const Part& P = parts[(size_t)g];
                if (pos[(size_t)g] >= P.n[(size_t)qq]) continue;
                const uint32_t o = score_to_order(P.sims[(size_t)qq * P.k + pos[(size_t)g]]);
                if (best < 0 || o > best_o) { best = g; best_o = o; }
            }
            if (best < 0) break;
            const Part& P = parts[(size_t)best];
            const size_t src = (size_t)qq * P.k + pos[(size_t)best];
            const size_t dst = (size_t)qq * topk + filled;
            if (out_ids) out_ids[dst] = P.ids[src];
            if (out_sims) out_sims[dst] = P.sims[src];
            if (out_keys) memcpy(out_keys + dst * dim, P.keys.data() + src * dim, (size_t)dim * 4);
            pos[(size_t)best]++;
            filled++;
        }
        if (out_n) out_n[qq] = filled;
    }
    return 0;
}

static int multi_find_batch(tfp_store* top, const float* q, int nq, int dim, int topk, int* out_n, int64_t* out_ids,
                            float* out_sims, float* out_keys, int* out_path) {
    tfp_multi* m = top->multi;
    std::shared_lock<RwMutex> lk(top->rw);
    if (dim != top->dim)                                                             // store.go:480-482 (fires on -1 too)
        return fail(1, "Try to find key with length %d when existing length is %d", dim, top->dim);
    if (topk < 1) return fail(1, "opts.TopK = %d, must be >= 1", topk);             // store.go:484-486
    if (!q || nq <= 0) return fail(2, "tfp_store_find: q is NULL or nq <= 0");
    const int G = (int)m->shards.size();
    std::vector<QInfo> qi((size_t)nq);
    for (int i = 0; i < nq; i++) {
        const float* qq = q + (size_t)i * dim;
        qi[(size_t)i].normalized = host_is_normalized(qq, dim) ? 1 : 0;
        qi[(size_t)i].mag1 = host_query_mag(qq, dim);
        qi[(size_t)i].pad = 0;
    }
    const size_t need = out_keys ? tfp_p2p_slot_bytes_with_keys(nq, topk, dim) : record_bytes(topk) * (size_t)nq;
    const bool on_device = m->p2p && (int64_t)G * topk <= 16384 && need <= m->slot_bytes;
    const int lane = multi_lane_acquire(m);
    MultiLane& L = m->lanes[(size_t)lane];
    int rc = on_device ? multi_find_device(m, L, q, nq, dim, topk, qi.data(), out_n, out_ids, out_sims, out_keys, out_path)
                       : multi_find_general(m, L, q, nq, dim, topk, qi.data(), out_n, out_ids, out_sims, out_keys, out_path);
    multi_lane_release(m, lane);
    return rc;
}

// Snapshot of a multi-GPU store: one file per shard (`path.g<i>`) and a manifest at `path` written LAST, which
// names the shard count and every shard's row count and id counter; a load checks the shard files against it.
struct MultiManifest {
    char magic[8];        // "TFPMULT1"
    int32_t shards, dim, normalized, reserved;
    int64_t rows[kMultiMaxShards], next_id[kMultiMaxShards];
    uint64_t tag;         // the host's generation stamp (see SnapHeader)
};

static int multi_save(tfp_store* top, const char* path, uint64_t tag) {
    tfp_multi* m = top->multi;
    std::lock_guard<std::mutex> mut(top->mut_mu);
    MultiManifest mf;
    memset(&mf, 0, sizeof(mf));
    memcpy(mf.magic, "TFPMULT1", 8);
    mf.shards = (int32_t)m->shards.size(); mf.dim = top->dim; mf.normalized = top->normalized ? 1 : 0; mf.tag = tag;
    int rc = run_shards(m, all_shards(m), [&](int g) -> int {
        std::string p = std::string(path) + ".g" + std::to_string(g);
        return store_save(m->shards[(size_t)g], p.c_str(), tag);
    });
    if (rc) return rc;
    for (size_t g = 0; g < m->shards.size(); g++) { mf.rows[g] = m->shards[g]->n; mf.next_id[g] = m->shards[g]->next_id; }
    return write_file_atomically(path, [&](FILE* f) { return fwrite(&mf, sizeof(mf), 1, f) == 1; });
}

static int multi_load(tfp_store* top, const char* path) {
    tfp_multi* m = top->multi;
    std::lock_guard<std::mutex> mut(top->mut_mu);
    std::unique_lock<RwMutex> lk(top->rw);
    if (top->dim != -1) return fail(2, "tfp_store_load: the store is not empty");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(2, "tfp_store_load: cannot open %s", path);
    MultiManifest mf;
    const bool ok = fread(&mf, sizeof(mf), 1, f) == 1 && memcmp(mf.magic, "TFPMULT1", 8) == 0;
    fclose(f);
    if (!ok) return fail(2, "tfp_store_load: %s is not a multi-GPU store snapshot", path);
    if (mf.shards != (int32_t)m->shards.size())
        return fail(2, "tfp_store_load: the snapshot has %d shards, this store has %d", (int)mf.shards, (int)m->shards.size());
    int rc = run_shards(m, all_shards(m), [&](int g) -> int {
        std::string p = std::string(path) + ".g" + std::to_string(g);
        int r = store_load(m->shards[(size_t)g], p.c_str());
        if (r) return r;
        tfp_store* sh = m->shards[(size_t)g];
        if (sh->n != mf.rows[g] || sh->next_id != mf.next_id[g])
            return fail(2, "tfp_store_load: %s does not belong to this snapshot (rows %lld vs %lld)", p.c_str(), (long long)sh->n,
                        (long long)mf.rows[g]);
        return 0;
    });
    if (rc) return rc;    // (shards that did load keep their rows; the host discards a store whose load failed)
    top->dim = mf.dim;
    top->normalized = mf.normalized != 0;
    for (tfp_store* sh : m->shards) {
        std::unique_lock<RwMutex> l2(sh->rw);
        if (sh->dim == -1) sh->dim = mf.dim;
        if (!top->normalized) sh->normalized = false;
    }
    return 0;
}
