// block_manager.cpp -- host side of libb200attn.so: error text + the KV page allocator
// (SURVEY.md section 8(a) row a9, section 8(f) rank 1).
//
// The reference never allocates KV pages itself; it passes the knobs that size the pool to the
// engine it launches: block_size (reference benchmark-vllm/models.py:10), max_num_seqs (:9),
// max_model_len (:15), gpu_memory_utilization (:8).  This is the allocator an engine behind
// that contract needs every decode step: a LIFO free list of physical blocks and, per sequence,
// the ordered list of blocks it owns.  It emits exactly the three arrays the kernels consume:
// block_tables[B, max_nb], seq_lens[B], slot_mapping[T].
//
// Plain C++17, no CUDA: all outputs are host arrays the caller copies (pinned -> device).
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <unordered_map>
#include <vector>

#include "b200attn.h"

// ---------------------------------------------------------------------------------------------
// error text
// ---------------------------------------------------------------------------------------------
static thread_local char g_b2_err[512] = "";

void b2_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_b2_err, sizeof(g_b2_err), fmt, ap);
    va_end(ap);
}

extern "C" int b2_version(void) { return B2_ABI_VERSION; }
extern "C" const char* b2_last_error(void) { return g_b2_err; }

// ---------------------------------------------------------------------------------------------
// block manager
// ---------------------------------------------------------------------------------------------
struct b2_seq {
    std::vector<int32_t> blocks;  // physical ids, logical order
    int64_t len = 0;              // tokens written so far
};

struct b2_block_manager {
    int64_t num_blocks;
    int bs;
    std::vector<int32_t> free_list;  // stack: back() is handed out next
    std::unordered_map<int64_t, b2_seq> seqs;
};

static inline int64_t blocks_for(int64_t tokens, int bs) { return (tokens + bs - 1) / bs; }

extern "C" b2_block_manager* b2_bm_create(int64_t num_blocks, int bs) {
    if (num_blocks <= 0 || num_blocks > INT32_MAX || bs <= 0) {
        b2_set_error("bm_create: bad args num_blocks=%lld bs=%d", (long long)num_blocks, bs);
        return nullptr;
    }
    b2_block_manager* bm = new (std::nothrow) b2_block_manager();
    if (!bm) {
        b2_set_error("bm_create: out of host memory");
        return nullptr;
    }
    bm->num_blocks = num_blocks;
    bm->bs = bs;
    bm->free_list.resize(num_blocks);
    // hand out low ids first
    for (int64_t i = 0; i < num_blocks; ++i) bm->free_list[i] = static_cast<int32_t>(num_blocks - 1 - i);
    return bm;
}

extern "C" void b2_bm_destroy(b2_block_manager* bm) { delete bm; }

extern "C" int64_t b2_bm_num_free(const b2_block_manager* bm) {
    return bm ? static_cast<int64_t>(bm->free_list.size()) : B2_EINVAL;
}

extern "C" int64_t b2_bm_num_seqs(const b2_block_manager* bm) {
    return bm ? static_cast<int64_t>(bm->seqs.size()) : B2_EINVAL;
}

extern "C" int64_t b2_bm_seq_len(const b2_block_manager* bm, int64_t seq_id) {
    if (!bm) return B2_EINVAL;
    auto it = bm->seqs.find(seq_id);
    if (it == bm->seqs.end()) return B2_ENOTFOUND;
    return it->second.len;
}

extern "C" int b2_bm_can_append(const b2_block_manager* bm, int64_t seq_id, int64_t n_tokens) {
    if (!bm || n_tokens < 0) return 0;
    auto it = bm->seqs.find(seq_id);
    int64_t have = 0, len = 0;
    if (it != bm->seqs.end()) {
        have = static_cast<int64_t>(it->second.blocks.size());
        len = it->second.len;
    }
    const int64_t need = blocks_for(len + n_tokens, bm->bs) - have;
    return need <= static_cast<int64_t>(bm->free_list.size()) ? 1 : 0;
}

static int grow_to(b2_block_manager* bm, b2_seq& s, int64_t tokens) {
    const int64_t want = blocks_for(tokens, bm->bs);
    const int64_t have = static_cast<int64_t>(s.blocks.size());
    if (want <= have) return B2_OK;
    if (want - have > static_cast<int64_t>(bm->free_list.size())) {
        b2_set_error("block pool exhausted: need %lld more blocks, %zu free",
                     (long long)(want - have), bm->free_list.size());
        return B2_ENOMEM;  // nothing was taken: the call is all-or-nothing
    }
    for (int64_t i = have; i < want; ++i) {
        s.blocks.push_back(bm->free_list.back());
        bm->free_list.pop_back();
    }
    return B2_OK;
}

extern "C" int b2_bm_alloc_seq(b2_block_manager* bm, int64_t seq_id, int64_t n_tokens) {
    if (!bm || n_tokens < 0) {
        b2_set_error("bm_alloc_seq: bad args");
        return B2_EINVAL;
    }
    if (bm->seqs.count(seq_id)) {
        b2_set_error("bm_alloc_seq: sequence %lld already exists", (long long)seq_id);
        return B2_EEXISTS;
    }
    b2_seq s;
    const int rc = grow_to(bm, s, n_tokens);
    if (rc != B2_OK) return rc;
    bm->seqs.emplace(seq_id, std::move(s));
    return B2_OK;
}

extern "C" int b2_bm_append(b2_block_manager* bm, int64_t seq_id, int64_t n_new, int64_t* slots_out) {
    if (!bm || n_new < 0 || (n_new > 0 && !slots_out)) {
        b2_set_error("bm_append: bad args");
        return B2_EINVAL;
    }
    auto it = bm->seqs.find(seq_id);
    if (it == bm->seqs.end()) {
        b2_set_error("bm_append: unknown sequence %lld", (long long)seq_id);
        return B2_ENOTFOUND;
    }
    b2_seq& s = it->second;
    const int rc = grow_to(bm, s, s.len + n_new);
    if (rc != B2_OK) return rc;
    const int bs = bm->bs;
    for (int64_t i = 0; i < n_new; ++i) {
        const int64_t pos = s.len + i;
        slots_out[i] = static_cast<int64_t>(s.blocks[pos / bs]) * bs + pos % bs;
    }
    s.len += n_new;
    return B2_OK;
}

extern "C" int b2_bm_free_seq(b2_block_manager* bm, int64_t seq_id) {
    if (!bm) return B2_EINVAL;
    auto it = bm->seqs.find(seq_id);
    if (it == bm->seqs.end()) {
        b2_set_error("bm_free_seq: unknown sequence %lld", (long long)seq_id);
        return B2_ENOTFOUND;
    }
    // return in reverse so the most recently used block is reused first (L2-warm)
    for (auto rit = it->second.blocks.rbegin(); rit != it->second.blocks.rend(); ++rit)
        bm->free_list.push_back(*rit);
    bm->seqs.erase(it);
    return B2_OK;
}

extern "C" int b2_bm_fill_tables(const b2_block_manager* bm, const int64_t* seq_ids, int B,
                                 int32_t* block_tables_out, int max_nb, int32_t* seq_lens_out) {
    if (!bm || B < 0 || max_nb <= 0 || (B > 0 && (!seq_ids || !block_tables_out || !seq_lens_out))) {
        b2_set_error("bm_fill_tables: bad args");
        return B2_EINVAL;
    }
    for (int b = 0; b < B; ++b) {
        auto it = bm->seqs.find(seq_ids[b]);
        if (it == bm->seqs.end()) {
            b2_set_error("bm_fill_tables: unknown sequence %lld", (long long)seq_ids[b]);
            return B2_ENOTFOUND;
        }
        const b2_seq& s = it->second;
        const int64_t used = blocks_for(s.len, bm->bs);
        if (used > max_nb) {
            b2_set_error("bm_fill_tables: sequence %lld needs %lld blocks > max_nb %d",
                         (long long)seq_ids[b], (long long)used, max_nb);
            return B2_EINVAL;
        }
        int32_t* row = block_tables_out + static_cast<int64_t>(b) * max_nb;
        if (used) memcpy(row, s.blocks.data(), used * sizeof(int32_t));
        if (used < max_nb) memset(row + used, 0, (max_nb - used) * sizeof(int32_t));
        seq_lens_out[b] = static_cast<int32_t>(s.len);
    }
    return B2_OK;
}

extern "C" int b2_bm_plan_decode(b2_block_manager* bm, const int64_t* seq_ids, int B,
                                 int32_t* block_tables_out, int max_nb, int32_t* seq_lens_out,
                                 int64_t* slot_mapping_out) {
    if (!bm || B < 0 || (B > 0 && (!seq_ids || !slot_mapping_out))) {
        b2_set_error("bm_plan_decode: bad args");
        return B2_EINVAL;
    }
    // all-or-nothing: count the new blocks the whole batch needs before touching anything
    int64_t need = 0;
    for (int b = 0; b < B; ++b) {
        auto it = bm->seqs.find(seq_ids[b]);
        if (it == bm->seqs.end()) {
            b2_set_error("bm_plan_decode: unknown sequence %lld", (long long)seq_ids[b]);
            return B2_ENOTFOUND;
        }
        const b2_seq& s = it->second;
        if (blocks_for(s.len + 1, bm->bs) > max_nb) {
            b2_set_error("bm_plan_decode: sequence %lld would exceed max_nb %d",
                         (long long)seq_ids[b], max_nb);
            return B2_EINVAL;
        }
        need += blocks_for(s.len + 1, bm->bs) - static_cast<int64_t>(s.blocks.size()) > 0
                    ? blocks_for(s.len + 1, bm->bs) - static_cast<int64_t>(s.blocks.size())
                    : 0;
    }
    if (need > static_cast<int64_t>(bm->free_list.size())) {
        b2_set_error("block pool exhausted: decode step needs %lld blocks, %zu free",
                     (long long)need, bm->free_list.size());
        return B2_ENOMEM;
    }
    for (int b = 0; b < B; ++b) {
        const int rc = b2_bm_append(bm, seq_ids[b], 1, slot_mapping_out + b);
        if (rc != B2_OK) return rc;
    }
    return b2_bm_fill_tables(bm, seq_ids, B, block_tables_out, max_nb, seq_lens_out);
}
