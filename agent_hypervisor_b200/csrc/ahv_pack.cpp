// ahv_pack.cpp - host-side canonical packer: the bytes SemanticDelta.compute_hash hashes
// (audit/delta.py:43-62: json.dumps({...7 keys...}, sort_keys=True).encode()) written
// straight into the arena from columnar inputs, multi-threaded.  Host code by design
// (BASELINE.json north_star item 1: the host packs, the GPU hashes); SURVEY 8f rank 1.
// The emitter itself (escaping, isoformat, key order) lives in pack_core.cuh, shared with the device packer.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include <unistd.h>

#include "../../include/ahv.h"

namespace ahv {
void set_error(const char *fmt, ...);
}

#include "pack_core.cuh"

using namespace ahv_pack;

namespace {

// Worker threads that stay around between calls (a replay packs chunk after chunk; starting and joining 16
// threads per pass costs about as much as packing 50 000 deltas).  One job at a time; the pool is rebuilt
// in a forked child, whose copy has no threads.
class Pool {
public:
    // One pool per process.  A forked child gets a NEW one: the parent's threads do not exist there and the
    // copied mutexes / condition variables may hold waiter state, so nothing of the old object is touched
    // again (it is left behind, like the pool itself at exit: the workers just end with the process).
    static Pool &get()
    {
        static std::mutex guard;
        static Pool *pool = nullptr;
        static pid_t owner = 0;
        std::lock_guard<std::mutex> lk(guard);
        if (!pool || owner != getpid()) {
            pool = new Pool();
            owner = getpid();
        }
        return *pool;
    }
    // f(lo, hi) over [0, n) in blocks handed out on demand (a worker that loses its core to another thread
    // of the process then only delays one block, not a fixed share); returns when every block is done
    void run(uint64_t n, const std::function<void(uint64_t, uint64_t)> &f)
    {
        std::lock_guard<std::mutex> one_job(job_mutex_);
        ensure_threads();
        const unsigned t = (unsigned)workers_.size() + 1;                 // the caller works too
        if (n < 4096 || t == 1) {
            f(0, n);
            return;
        }
        {
            std::lock_guard<std::mutex> lk(m_);
            f_ = &f;
            n_ = n;
            grain_ = std::max<uint64_t>(1024, n / (8ull * t));
            next_.store(0, std::memory_order_relaxed);
            pending_ = t - 1;
            ++generation_;
        }
        cv_.notify_all();
        blocks(n, grain_, f);
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [&] { return pending_ == 0; });
        f_ = nullptr;
    }

private:
    void blocks(uint64_t n, uint64_t grain, const std::function<void(uint64_t, uint64_t)> &f)
    {
        for (;;) {
            const uint64_t lo = next_.fetch_add(grain, std::memory_order_relaxed);
            if (lo >= n) return;
            f(lo, std::min<uint64_t>(n, lo + grain));
        }
    }
    void ensure_threads()
    {
        if (started_) return;
        started_ = true;
        unsigned t = std::thread::hardware_concurrency();
        if (t == 0) t = 1;
        t = std::min(t, 64u);
        const uint64_t gen = generation_;
        for (unsigned k = 0; k + 1 < t; ++k) workers_.emplace_back([this, k, gen] { loop(k, gen); });
    }
    void loop(unsigned /*k*/, uint64_t seen)
    {
        for (;;) {
            const std::function<void(uint64_t, uint64_t)> *f;
            uint64_t n, grain;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return stop_ || generation_ != seen; });
                if (stop_) return;
                seen = generation_;
                f = f_;
                n = n_;
                grain = grain_;
            }
            blocks(n, grain, *f);
            std::lock_guard<std::mutex> lk(m_);
            if (--pending_ == 0) done_.notify_one();
        }
    }
    std::mutex job_mutex_, m_;
    std::condition_variable cv_, done_;
    std::vector<std::thread> workers_;
    const std::function<void(uint64_t, uint64_t)> *f_ = nullptr;
    uint64_t n_ = 0, grain_ = 0, generation_ = 0;
    std::atomic<uint64_t> next_{0};
    unsigned pending_ = 0;
    bool started_ = false, stop_ = false;
};

template <class F>
void parallel_for(uint64_t n, F f)
{
    Pool::get().run(n, std::function<void(uint64_t, uint64_t)>(f));
}

}  // namespace

extern "C" int ahv_pack_deltas(const ahv_delta_columns *cols, uint64_t *out_offsets, uint8_t *out_arena,
                               uint64_t arena_capacity, uint64_t *out_bytes, uint32_t *out_parent_pos, uint32_t flags)
{
    if (!cols || !out_offsets || !out_bytes) {
        ahv::set_error("invalid argument: cols / out_offsets / out_bytes is null");
        return AHV_E_INVALID;
    }
    Cols c;
    c.n = cols->n;
    c.delta_id = {cols->delta_id_blob, cols->delta_id_off, nullptr};
    c.session_id = {cols->session_id_blob, cols->session_id_off, nullptr};
    c.agent_did = {cols->agent_did_blob, cols->agent_did_off, nullptr};
    c.turn_id = cols->turn_id;
    c.ts_us = cols->timestamp_us;
    c.ts_off = cols->timestamp_utcoffset_s;
    c.parent_kind = cols->parent_kind;
    c.parent_digest = cols->parent_digest;
    c.parent_other = {cols->parent_other_blob, cols->parent_other_off, nullptr};
    c.change_rowptr = cols->change_rowptr;
    c.ch_path = {cols->change_path_blob, cols->change_path_off, nullptr};
    c.ch_op = {cols->change_op_blob, cols->change_op_off, nullptr};
    c.ch_content = {cols->change_content_blob, cols->change_content_off, cols->change_content_null};
    c.ch_prev = {cols->change_prev_blob, cols->change_prev_off, cols->change_prev_null};
    const uint64_t n = c.n;
    // pass 1: sizes (out_offsets[i+1] = length of delta i), then an exclusive scan; skipped when the caller
    // says out_offsets still holds the result of its size query for these columns
    const bool have_offsets = (flags & AHV_PACK_OFFSETS_VALID) && out_arena && out_offsets[0] == 0;
    if (!have_offsets) {
        out_offsets[0] = 0;
        parallel_for(n, [&](uint64_t lo, uint64_t hi) {
            for (uint64_t i = lo; i < hi; ++i) {
                SinkT<false> s(nullptr);
                const uint64_t pp = emit_delta(s, c, i);
                out_offsets[i + 1] = s.n;
                // a position past 4 GiB cannot be stored: UINT32_MAX - 1 is never a valid one, so it reads as "other"
                if (out_parent_pos) out_parent_pos[i] = pp == UINT32_MAX ? UINT32_MAX : (uint32_t)std::min<uint64_t>(pp, UINT32_MAX - 1);
            }
        });
        for (uint64_t i = 0; i < n; ++i) out_offsets[i + 1] += out_offsets[i];
    }
    *out_bytes = out_offsets[n];
    if (!out_arena) return AHV_OK;                       // size query
    if (arena_capacity < out_offsets[n]) {
        ahv::set_error("arena capacity %llu < %llu bytes needed", (unsigned long long)arena_capacity,
                       (unsigned long long)out_offsets[n]);
        return AHV_E_TOOLARGE;
    }
    parallel_for(n, [&](uint64_t lo, uint64_t hi) {
        for (uint64_t i = lo; i < hi; ++i) {
            SinkT<true> s(out_arena + out_offsets[i]);
            emit_delta(s, c, i);
        }
    });
    return AHV_OK;
}
