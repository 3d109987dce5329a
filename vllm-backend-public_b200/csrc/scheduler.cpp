// scheduler.cpp -- continuous-batching step planner of libb200attn.so (host only).
// the caller of the hot path; it decides B and the raggedness every step.
//
// Limits come from the reference's launch contract: max_num_seqs (reference benchmark-vllm/models.py:9),
// max_model_len (:15), block_size (:10), the pool sized from gpu_memory_utilization (:8).  Policy and the
// request life cycle are documented in include/b200attn.h next to the b2_sched_* declarations.
//
// Built on the block manager's C API (block_manager.cpp); plain C++17, no CUDA.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <deque>
#include <unordered_map>
#include <vector>

#include "b200attn.h"

void b2_set_error(const char* fmt, ...);

namespace {

struct Req {
    int64_t id;
    int prompt_len;        // tokens to prefill (grows by the generated tokens when the request is recomputed)
    int orig_prompt;       // the request's own prompt length
    int max_new;           // sampling budget
    int generated = 0;     // tokens sampled so far
    int computed = 0;      // prompt tokens already in the cache
    int64_t admit_seq = 0; // admission order, for the preemption victim choice
    bool running = false;  // holds blocks
    bool resumed = false;  // being recomputed after a preemption: its next token is already known
};

}  // namespace

struct b2_scheduler {
    b2_block_manager* bm = nullptr;
    int64_t num_blocks = 0;
    int bs = 0, max_num_seqs = 0, max_model_len = 0, token_budget = 0, max_nb = 0;
    int64_t admit_counter = 0;
    std::unordered_map<int64_t, Req> reqs;
    std::deque<int64_t> waiting;       // FCFS; preempted requests re-enter at the front
    std::vector<int64_t> running;      // admission order; decode rows follow it
    std::vector<int64_t> finished_pending;   // finished by length after the previous plan
    // plan storage
    std::vector<int64_t> decode_ids, slot_mapping, prefill_ids, prefill_slots, preempted, finished;
    std::vector<int32_t> block_tables, seq_lens, prefill_start, prefill_len, prefill_bt;
    std::vector<uint8_t> prefill_last;
};

static void sched_release(b2_scheduler* s, Req& r) {
    if (r.running) {
        b2_bm_free_seq(s->bm, r.id);
        r.running = false;
        s->running.erase(std::remove(s->running.begin(), s->running.end(), r.id), s->running.end());
    }
}

extern "C" b2_scheduler* b2_sched_create(int64_t num_blocks, int bs, int max_num_seqs, int max_model_len,
                                         int max_num_batched_tokens) {
    if (num_blocks <= 0 || bs <= 0 || max_num_seqs <= 0 || max_model_len <= 0 || max_num_batched_tokens <= 0) {
        b2_set_error("sched_create: bad args");
        return nullptr;
    }
    b2_scheduler* s = new (std::nothrow) b2_scheduler();
    if (!s) return nullptr;
    s->bm = b2_bm_create(num_blocks, bs);
    if (!s->bm) {
        delete s;
        return nullptr;
    }
    s->num_blocks = num_blocks;
    s->bs = bs;
    s->max_num_seqs = max_num_seqs;
    s->max_model_len = max_model_len;
    s->token_budget = max_num_batched_tokens;
    s->max_nb = (max_model_len + bs - 1) / bs;
    return s;
}

extern "C" void b2_sched_destroy(b2_scheduler* s) {
    if (!s) return;
    b2_bm_destroy(s->bm);
    delete s;
}

extern "C" int b2_sched_add_request(b2_scheduler* s, int64_t req_id, int prompt_len, int max_new_tokens) {
    if (!s || prompt_len <= 0 || max_new_tokens <= 0) {
        b2_set_error("sched_add_request: bad args");
        return B2_EINVAL;
    }
    if (s->reqs.count(req_id)) {
        b2_set_error("sched_add_request: request %lld already exists", (long long)req_id);
        return B2_EEXISTS;
    }
    if (prompt_len >= s->max_model_len) {
        b2_set_error("sched_add_request: prompt of %d tokens leaves no room under max_model_len %d", prompt_len,
                     s->max_model_len);
        return B2_EINVAL;
    }
    // a request that can never fit the pool would dead-lock the queue
    if (((int64_t)prompt_len + 1 + s->bs - 1) / s->bs > s->num_blocks) {
        b2_set_error("sched_add_request: a prompt of %d tokens cannot fit a pool of %lld blocks", prompt_len,
                     (long long)s->num_blocks);
        return B2_ENOMEM;
    }
    Req r;
    r.id = req_id;
    r.prompt_len = prompt_len;
    r.orig_prompt = prompt_len;
    // A sequence caches prompt + generated - 1 tokens.  One that would outgrow the WHOLE pool can never finish: alone
    // in the running set it would be preempted, recomputed and preempted again for ever (a --num-gpu-blocks-override
    // or a direct b2_sched_create can make the pool smaller than one max_model_len sequence).  Its budget is cut to
    // what the pool can hold; it then finishes by length like any other request.
    const int64_t pool_tokens = s->num_blocks * (int64_t)s->bs;
    const int64_t fit_new = pool_tokens - prompt_len + 1;          // >= 2: the prompt + 1 fits (checked above)
    r.max_new = (int)std::min<int64_t>(max_new_tokens, fit_new);
    s->reqs.emplace(req_id, r);
    s->waiting.push_back(req_id);
    return B2_OK;
}

extern "C" int b2_sched_finish(b2_scheduler* s, int64_t req_id) {
    if (!s) return B2_EINVAL;
    auto it = s->reqs.find(req_id);
    if (it == s->reqs.end()) {
        b2_set_error("sched_finish: unknown request %lld", (long long)req_id);
        return B2_ENOTFOUND;
    }
    sched_release(s, it->second);
    s->waiting.erase(std::remove(s->waiting.begin(), s->waiting.end(), req_id), s->waiting.end());
    s->finished_pending.erase(std::remove(s->finished_pending.begin(), s->finished_pending.end(), req_id),
                              s->finished_pending.end());
    s->reqs.erase(it);
    return B2_OK;
}

extern "C" int64_t b2_sched_num_waiting(const b2_scheduler* s) { return s ? (int64_t)s->waiting.size() : B2_EINVAL; }
extern "C" int64_t b2_sched_num_running(const b2_scheduler* s) { return s ? (int64_t)s->running.size() : B2_EINVAL; }
extern "C" int64_t b2_sched_num_free_blocks(const b2_scheduler* s) { return s ? b2_bm_num_free(s->bm) : B2_EINVAL; }
extern "C" int64_t b2_sched_generated(const b2_scheduler* s, int64_t req_id) {
    if (!s) return B2_EINVAL;
    auto it = s->reqs.find(req_id);
    return it == s->reqs.end() ? (int64_t)B2_ENOTFOUND : (int64_t)it->second.generated;
}

// ---------------------------------------------------------------------------------------------
// one step
// ---------------------------------------------------------------------------------------------
static inline int64_t blocks_for(int64_t tokens, int bs) { return (tokens + bs - 1) / bs; }

// Put a running request back at the head of the waiting queue; its KV is dropped and will be recomputed over
// prompt + all but the last sampled token (that one is written by its first decode step after the recompute).
static void sched_preempt(b2_scheduler* s, Req& r) {
    sched_release(s, r);
    if (r.generated > 0) {
        r.prompt_len = r.orig_prompt + r.generated - 1;       // tokens whose KV existed: P + g - 1
        r.resumed = true;
    }
    r.computed = 0;
    s->waiting.push_front(r.id);
    s->preempted.push_back(r.id);
}

extern "C" int b2_sched_step(b2_scheduler* s, b2_step_plan* plan) {
    if (!s || !plan) {
        b2_set_error("sched_step: bad args");
        return B2_EINVAL;
    }
    const int bs = s->bs, max_nb = s->max_nb;
    s->decode_ids.clear(); s->slot_mapping.clear(); s->block_tables.clear(); s->seq_lens.clear();
    s->prefill_ids.clear(); s->prefill_slots.clear(); s->prefill_start.clear(); s->prefill_len.clear();
    s->prefill_last.clear(); s->prefill_bt.clear(); s->preempted.clear(); s->finished.clear();

    // 0. requests that finished by length when the previous plan ran: free them now
    for (int64_t id : s->finished_pending) {
        auto it = s->reqs.find(id);
        if (it == s->reqs.end()) continue;
        sched_release(s, it->second);
        s->reqs.erase(it);
        s->finished.push_back(id);
    }
    s->finished_pending.clear();

    // 1. decode: every running request whose prompt is complete writes one token.  Preempt from the back of the
    //    admission order until the step fits the pool.
    auto decodable = [&](const Req& r) { return r.running && r.computed == r.prompt_len && r.generated >= 1; };
    for (;;) {
        int64_t need = 0;
        for (int64_t id : s->running) {
            const Req& r = s->reqs.at(id);
            if (!decodable(r)) continue;
            const int64_t len = b2_bm_seq_len(s->bm, id);
            need += blocks_for(len + 1, bs) - blocks_for(len, bs);
        }
        if (need <= b2_bm_num_free(s->bm) || s->running.empty()) break;
        if (s->running.size() == 1) {
            // the only running sequence holds every used block and still cannot grow: preempting it would only
            // recompute it into the same wall.  End it by length (reported in finished_ids of this plan).
            Req& sole = s->reqs.at(s->running.back());
            const int64_t id = sole.id;
            sched_release(s, sole);
            s->reqs.erase(id);
            s->finished.push_back(id);
            break;
        }
        sched_preempt(s, s->reqs.at(s->running.back()));
    }
    for (int64_t id : s->running) {
        Req& r = s->reqs.at(id);
        if (!decodable(r)) continue;
        int64_t slot = 0;
        const int rc = b2_bm_append(s->bm, id, 1, &slot);
        if (rc != B2_OK) return rc;                      // cannot happen: the step was sized above
        s->decode_ids.push_back(id);
        s->slot_mapping.push_back(slot);
        r.generated += 1;                                // the token this step samples
        r.resumed = false;
        const int64_t len = b2_bm_seq_len(s->bm, id);
        // the sequence now holds len cached tokens + the one just sampled: stop at max_model_len tokens in total
        if (r.generated >= r.max_new || len + 1 >= s->max_model_len) s->finished_pending.push_back(id);
    }
    const int n_decode = (int)s->decode_ids.size();
    if (n_decode) {
        s->block_tables.resize((size_t)n_decode * max_nb);
        s->seq_lens.resize(n_decode);
        const int rc = b2_bm_fill_tables(s->bm, s->decode_ids.data(), n_decode, s->block_tables.data(), max_nb,
                                         s->seq_lens.data());
        if (rc != B2_OK) return rc;
    }

    // 2. prefill within the remaining token budget: unfinished prompts of running requests first (admission
    //    order), then new admissions, first come first served.  Never preempts; a chunk shrinks to the free blocks.
    int budget = s->token_budget - n_decode;
    auto plan_chunk = [&](Req& r) -> bool {             // false: nothing more fits this step
        if (budget <= 0) return false;
        const int remaining = r.prompt_len - r.computed;
        int chunk = std::min(remaining, budget);
        const int64_t room = b2_bm_num_free(s->bm) * bs + (blocks_for(r.computed, bs) * bs - r.computed);
        if (chunk > room) chunk = (int)room;
        if (chunk <= 0) return false;
        if (!r.running) {
            if (b2_bm_alloc_seq(s->bm, r.id, 0) != B2_OK) return false;
            r.running = true;
            r.admit_seq = s->admit_counter++;
            s->running.push_back(r.id);
        }
        const size_t off = s->prefill_slots.size();
        s->prefill_slots.resize(off + chunk);
        if (b2_bm_append(s->bm, r.id, chunk, s->prefill_slots.data() + off) != B2_OK) return false;
        s->prefill_ids.push_back(r.id);
        s->prefill_start.push_back(r.computed);
        s->prefill_len.push_back(chunk);
        r.computed += chunk;
        budget -= chunk;
        uint8_t last = 0;
        if (r.computed == r.prompt_len) {
            if (r.resumed) {
                last = 2;                                // recompute done: the next token is already known
            } else {
                last = 1;
                r.generated = 1;                         // token #1 is sampled from this chunk's last position
                if (r.generated >= r.max_new || r.prompt_len + 1 >= s->max_model_len) s->finished_pending.push_back(r.id);
            }
        }
        s->prefill_last.push_back(last);
        return true;
    };
    for (size_t i = 0; i < s->running.size() && budget > 0; ++i) {
        Req& r = s->reqs.at(s->running[i]);
        if (r.computed < r.prompt_len) plan_chunk(r);
    }
    // A step that had to preempt admits nothing (vLLM's rule): the victim sits at the head of the queue and would take the
    // blocks just freed, shrink its chunk to them and be evicted again at the next block boundary.
    const bool admit = s->preempted.empty();
    const int64_t watermark = std::max<int64_t>(1, s->num_blocks / 100);
    while (admit && budget > 0 && !s->waiting.empty() && (int)s->running.size() < s->max_num_seqs) {
        Req& r = s->reqs.at(s->waiting.front());
        // do not start a prompt whose first chunk cannot take at least one block's worth (or all of it)
        if (b2_bm_num_free(s->bm) == 0) break;
        // a preempted request comes back only when its whole recompute fits, with room left for the running
        // sequences' next block crossings -- otherwise it is admitted on the freed blocks and evicted again
        if (r.resumed) {
            const int64_t want = blocks_for((int64_t)r.prompt_len + 1, bs) + (s->running.empty() ? 0 : watermark);
            if (b2_bm_num_free(s->bm) < want) break;
        }
        s->waiting.pop_front();
        if (!plan_chunk(r)) {
            s->waiting.push_front(r.id);
            break;
        }
    }
    const int n_prefill = (int)s->prefill_ids.size();
    if (n_prefill) {
        s->prefill_bt.resize((size_t)n_prefill * max_nb);
        std::vector<int32_t> lens(n_prefill);
        const int rc = b2_bm_fill_tables(s->bm, s->prefill_ids.data(), n_prefill, s->prefill_bt.data(), max_nb, lens.data());
        if (rc != B2_OK) return rc;
    }

    plan->n_decode = n_decode;
    plan->max_nb = max_nb;
    plan->decode_ids = s->decode_ids.data();
    plan->block_tables = s->block_tables.data();
    plan->seq_lens = s->seq_lens.data();
    plan->slot_mapping = s->slot_mapping.data();
    plan->n_prefill = n_prefill;
    plan->n_prefill_tokens = (int)s->prefill_slots.size();
    plan->prefill_ids = s->prefill_ids.data();
    plan->prefill_start = s->prefill_start.data();
    plan->prefill_len = s->prefill_len.data();
    plan->prefill_last = s->prefill_last.data();
    plan->prefill_slots = s->prefill_slots.data();
    plan->prefill_block_tables = s->prefill_bt.data();
    plan->n_preempted = (int)s->preempted.size();
    plan->preempted_ids = s->preempted.data();
    plan->n_finished = (int)s->finished.size();
    plan->finished_ids = s->finished.data();
    return B2_OK;
}
