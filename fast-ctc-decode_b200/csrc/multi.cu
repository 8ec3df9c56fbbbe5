// multi.cu — one call, several GPUs: static read-index sharding with one host thread per device.
//
// The reference decodes one read per call and releases the GIL (src/lib.rs:199) so that callers spread reads
// over Python threads.  The batched equivalent of that affordance: the reads of one host array are cut into
// contiguous ranges [g*N/G, (g+1)*N/G), every range is decoded by its own host thread through its own handle
// (stream, staging and scratch of one device), and the results land in the caller's single set of host arrays
// in read order.  Reads are independent, so no data moves between devices and no collective is involved.
#include <cuda_runtime.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/fctc.h"

namespace {

struct DeviceSlot {
    fctc_handle* h = nullptr;
    std::mutex busy;  // one shard at a time per handle (a device listed twice runs its shards one after the other)
};

std::mutex g_table_lock;
std::map<int, DeviceSlot*> g_table;
thread_local std::string g_multi_error;

DeviceSlot* slot_of(int device, std::string* err) {
    std::lock_guard<std::mutex> lk(g_table_lock);
    auto it = g_table.find(device);
    if (it != g_table.end()) return it->second;
    fctc_handle* h = nullptr;
    const int rc = fctc_create(device, &h);
    if (rc != 0) {
        *err = std::string("device ") + std::to_string(device) + ": " + fctc_last_error(nullptr);
        return nullptr;
    }
    DeviceSlot* s = new DeviceSlot();
    s->h = h;
    g_table[device] = s;
    return s;
}

inline int64_t roundup32(int64_t v) { return (v + 31) & ~int64_t(31); }

// Runs shard(g, lo, hi, handle) for every device entry on its own thread; returns the first failure.
template <class Shard>
int run_shards(const int32_t* devices, int32_t n_devices, int64_t n_reads, Shard&& shard) {
    if (!devices || n_devices <= 0) {
        g_multi_error = "devices must list at least one device ordinal";
        return FCTC_E_ARGUMENT;
    }
    std::vector<DeviceSlot*> slots(n_devices, nullptr);
    for (int g = 0; g < n_devices; ++g) {
        std::string err;
        slots[g] = slot_of(devices[g], &err);
        if (!slots[g]) {
            g_multi_error = err;
            return FCTC_E_CUDA;
        }
    }
    std::vector<int> rc(n_devices, 0);
    std::vector<std::string> msg(n_devices);
    std::vector<std::thread> threads;
    threads.reserve(n_devices);
    for (int g = 0; g < n_devices; ++g) {
        threads.emplace_back([&, g]() {
            const int64_t lo = n_reads * g / n_devices, hi = n_reads * (g + 1) / n_devices;
            if (hi <= lo) return;
            std::lock_guard<std::mutex> lk(slots[g]->busy);
            rc[g] = shard(g, lo, hi, slots[g]->h);
            if (rc[g] < 0) msg[g] = fctc_last_error(slots[g]->h);
        });
    }
    for (std::thread& t : threads) t.join();
    for (int g = 0; g < n_devices; ++g)
        if (rc[g] < 0) {
            g_multi_error = "device " + std::to_string(devices[g]) + " (reads " + std::to_string(n_reads * g / n_devices) + ".." +
                            std::to_string(n_reads * (g + 1) / n_devices) + "): " + msg[g];
            return rc[g];
        }
    return 0;
}

}  // namespace

extern "C" const char* fctc_multi_last_error(void) { return g_multi_error.c_str(); }

extern "C" void fctc_multi_release(void) {
    std::lock_guard<std::mutex> lk(g_table_lock);
    for (auto& kv : g_table) {
        std::lock_guard<std::mutex> busy(kv.second->busy);
        fctc_destroy(kv.second->h);
        kv.second->h = nullptr;
    }
    for (auto& kv : g_table) delete kv.second;
    g_table.clear();
}

extern "C" int fctc_viterbi_batch_multi(const int32_t* devices, int32_t n_devices, const float* post_host, int64_t n_reads,
                                        int64_t n_rows, int32_t n_labels, int32_t qstring, float qscale, float qbias,
                                        int32_t collapse_repeats, uint8_t* labels, void* quals, int32_t qual_bytes,
                                        int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                                        int32_t* status, int64_t* pool_used) {
    g_multi_error.clear();
    if (pool_used) *pool_used = 0;
    if (n_reads < 0 || n_rows <= 0 || n_labels <= 0) {
        g_multi_error = "n_reads >= 0, n_rows >= 1 and n_labels >= 1 are required";
        return FCTC_E_ARGUMENT;
    }
    const int64_t stride = roundup32(n_rows);
    if (pool_capacity < n_reads * stride) {
        g_multi_error = "result pool too small: n_reads * round_up(n_rows, 32) entries are needed";
        return FCTC_E_CAPACITY;
    }
    if (qual_bytes != 1 && qual_bytes != 4) {
        g_multi_error = "qual_bytes must be 1 or 4";
        return FCTC_E_ARGUMENT;
    }
    const int64_t read_elems = n_rows * int64_t(n_labels);
    const int rc = run_shards(devices, n_devices, n_reads, [&](int, int64_t lo, int64_t hi, fctc_handle* h) -> int {
        // slot form: read n owns entries [n * stride, ...), so every shard writes straight into its part of the pool
        const int64_t base = lo * stride;
        int64_t used = 0;
        const int r = fctc_viterbi_batch(h, post_host + lo * read_elems, hi - lo, n_rows, n_labels, qstring, qscale, qbias,
                                         collapse_repeats, labels + base,
                                         quals ? static_cast<uint8_t*>(quals) + base * qual_bytes : nullptr, qual_bytes,
                                         path + base, (hi - lo) * stride, offsets + lo, lens + lo, status + lo, &used);
        if (r < 0) return r;
        for (int64_t n = lo; n < hi; ++n) offsets[n] += base;
        return 0;
    });
    if (rc == 0 && pool_used) *pool_used = n_reads * stride;
    return rc;
}

extern "C" int fctc_beam_search_batch_multi(const int32_t* devices, int32_t n_devices, const float* post_host,
                                            int64_t n_reads, int64_t n_rows, int32_t n_labels, int32_t beam_size,
                                            float beam_cut_threshold, int32_t collapse_repeats, uint8_t* labels,
                                            int32_t* path, int64_t pool_capacity, int64_t* offsets, int32_t* lens,
                                            int32_t* status, int64_t* pool_used) {
    g_multi_error.clear();
    if (pool_used) *pool_used = 0;
    if (n_reads < 0 || n_rows < 0 || n_labels <= 0) {
        g_multi_error = "n_reads >= 0, n_rows >= 0 and n_labels >= 1 are required";
        return FCTC_E_ARGUMENT;
    }
    const int64_t stride = std::max<int64_t>(roundup32(n_rows), 32);
    if (pool_capacity < n_reads * stride) {
        g_multi_error = "result pool too small: n_reads * round_up(n_rows, 32) entries are needed";
        return FCTC_E_CAPACITY;
    }
    const int64_t read_elems = n_rows * int64_t(n_labels);
    std::vector<int64_t> used(std::max(n_devices, 1), 0);
    const int rc = run_shards(devices, n_devices, n_reads, [&](int g, int64_t lo, int64_t hi, fctc_handle* h) -> int {
        // packed pools: shard g fills the pool from entry lo * stride on (room for all of its reads), its offsets are
        // moved there; shards are in read order, reads inside a shard wherever the kernel's scheduling put them
        const int64_t base = lo * stride;
        const int r = fctc_beam_search_batch(h, post_host + lo * read_elems, hi - lo, n_rows, n_labels, beam_size,
                                             beam_cut_threshold, collapse_repeats, labels + base, path + base,
                                             (hi - lo) * stride, offsets + lo, lens + lo, status + lo, &used[g]);
        if (r < 0) return r;
        for (int64_t n = lo; n < hi; ++n) offsets[n] += base;
        return 0;
    });
    if (rc == 0 && pool_used) {
        int64_t total = 0;
        for (int g = 0; g < n_devices; ++g) total += used[g];
        *pool_used = total;
    }
    return rc;
}
