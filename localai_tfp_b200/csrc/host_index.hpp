// host_index.hpp -- exact-match index of the store: row hash -> slot(s).
//
// Replaces the reference's "keys are kept sorted, look up by binary search"
// (backend/go/stores/store.go:60-64 findInSortedSlice, used by StoresDelete
// :226 and StoresGet :286, and the merge in StoresSet :150-189).  Hashes come
// from the device (ingest.cuh row_stats_kernel); the index itself never reads
// key bytes (the host only compares the keys of ONE request with each other, to
// put them in the reference's sorted order: tfp_store.cu sorted_order).
// Several slots may share a hash (true duplicates stored by one
// StoresSet call, or a 64-bit collision); callers confirm every candidate on
// the device before trusting it.
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

namespace tfp {

class FlatIndex {
    enum : uint8_t { kEmpty = 0, kFull = 1, kTomb = 2 };
    std::vector<uint64_t> hash_;
    std::vector<uint32_t> slot_;
    std::vector<uint8_t> state_;
    size_t mask_ = 0, full_ = 0, tombs_ = 0;

    static size_t start(uint64_t h, size_t mask) { return (size_t)(h ^ (h >> 29)) & mask; }

    void rehash(size_t want) {
        size_t cap = 1024;
        while (cap < want * 2) cap <<= 1;
        std::vector<uint64_t> oh;
        std::vector<uint32_t> os;
        std::vector<uint8_t> ost;
        oh.swap(hash_); os.swap(slot_); ost.swap(state_);
        hash_.assign(cap, 0); slot_.assign(cap, 0); state_.assign(cap, kEmpty);
        mask_ = cap - 1; full_ = 0; tombs_ = 0;
        for (size_t i = 0; i < ost.size(); i++)
            if (ost[i] == kFull) insert(oh[i], os[i]);
    }

public:
    FlatIndex() { rehash(512); }
    size_t size() const { return full_; }
    void reserve(size_t n) { if ((n + tombs_) * 10 > (mask_ + 1) * 6) rehash(n); }

    void insert(uint64_t h, uint32_t slot) {
        if ((full_ + tombs_ + 1) * 10 > (mask_ + 1) * 6) rehash(full_ + 1);
        size_t i = start(h, mask_);
        while (state_[i] == kFull) i = (i + 1) & mask_;
        if (state_[i] == kTomb) tombs_--;
        hash_[i] = h; slot_[i] = slot; state_[i] = kFull; full_++;
    }
    // Every stored slot whose hash equals h, in probe order.
    template <class F> void for_each(uint64_t h, F&& f) const {
        size_t i = start(h, mask_);
        while (state_[i] != kEmpty) {
            if (state_[i] == kFull && hash_[i] == h) f(slot_[i]);
            i = (i + 1) & mask_;
        }
    }
    bool erase(uint64_t h, uint32_t slot) {
        size_t i = start(h, mask_);
        while (state_[i] != kEmpty) {
            if (state_[i] == kFull && hash_[i] == h && slot_[i] == slot) {
                state_[i] = kTomb; full_--; tombs_++;
                return true;
            }
            i = (i + 1) & mask_;
        }
        return false;
    }
    bool relocate(uint64_t h, uint32_t from, uint32_t to) {
        size_t i = start(h, mask_);
        while (state_[i] != kEmpty) {
            if (state_[i] == kFull && hash_[i] == h && slot_[i] == from) { slot_[i] = to; return true; }
            i = (i + 1) & mask_;
        }
        return false;
    }
};

}  // namespace tfp
