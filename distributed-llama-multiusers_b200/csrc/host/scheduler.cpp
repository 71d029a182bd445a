// Continuous-batching request scheduler (the intended behaviour of the fork's Request /
// RequestQueue / inference_loop: src/Request.hpp:21-64, src/app.cpp:314-402 -- minus its defects
// D1-D4: here every request keeps its own prompt, position and KV slot, prompts are prefilled in
// chunks, and any number of HTTP threads can submit concurrently).
//
// Policy per step: every running (decoding) request contributes one row; the remaining rows (at least a quarter
// of the batch while prompts are pending) go to requests that still have prompt tokens to evaluate, in admission
// order, at most prefill_chunk rows each.  Prompt rows that need no logits come FIRST; the last row of every prompt
// that completes in this step is moved in front of the decode rows, so that the rows needing logits form the
// contiguous tail [logits_from, n).
// Thread-safe: submit/cancel from any thread; next_batch/commit from the engine loop thread.
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace {

enum State { WAITING = 0, PREFILL = 1, DECODE = 2, DONE = 3 };

struct Req {
    long long id;
    std::vector<int> prompt;
    int max_new;
    int state = WAITING;
    int slot = -1;
    int n_prefilled = 0;     // prompt tokens already evaluated
    int pos = 0;             // next position to write
    int last_token = -1;     // token to feed on the next decode row
    int n_generated = 0;
    int pages = 0;           // KV pages reserved for this request while it holds a slot
    bool cancelled = false;
};

struct Sched {
    int max_rows, n_slots, prefill_chunk, max_seq_len;
    // KV page budget: a request is admitted only if the pages of its whole life (prompt + max_new tokens) fit next to
    // the reservations of the running ones, so the engine's page allocator can never run dry in the middle of a step.
    // 0 pages = no page accounting (admission by free slot only).
    int n_pages = 0, page_size = 32, pages_reserved = 0;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<long long> waiting;                       // FIFO admission
    std::vector<long long> running;                      // PREFILL/DECODE, admission order
    std::unordered_map<long long, Req> reqs;
    std::vector<int> free_slots;
    std::vector<int> released;                           // slots whose KV pages the engine should free
    size_t rr = 0;                                       // round-robin cursor when running > max_rows
    // rows of the batch handed out by next_batch (to validate commit)
    std::vector<long long> batch_sampling;
};

}  // namespace

extern "C" {

void* dl_sched_new(int max_rows, int n_slots, int prefill_chunk, int max_seq_len) {
    auto* s = new Sched();
    s->max_rows = max_rows;
    s->n_slots = n_slots;
    s->prefill_chunk = prefill_chunk > 0 ? prefill_chunk : max_rows;
    s->max_seq_len = max_seq_len;
    for (int i = n_slots - 1; i >= 0; i--) s->free_slots.push_back(i);
    return s;
}

void dl_sched_free(void* p) { delete (Sched*)p; }

// KV page pool of the engine (pages shared by all sequences) and tokens per page; call before the first submit
void dl_sched_set_pages(void* p, int n_pages, int page_size) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    s->n_pages = n_pages > 0 ? n_pages : 0;
    s->page_size = page_size > 0 ? page_size : 32;
}

static int pages_for(const Sched* s, int n_tokens) { return (n_tokens + s->page_size - 1) / s->page_size; }

// 0 = queued; -1 = prompt (+ max_new) does not fit max_seq_len; -2 = duplicate id; -3 = empty prompt
int dl_sched_submit(void* p, long long req_id, const int* tokens, int n, int max_new) {
    auto* s = (Sched*)p;
    if (n <= 0) return -3;
    if (n >= s->max_seq_len) return -1;
    std::lock_guard<std::mutex> g(s->mu);
    if (s->reqs.count(req_id)) return -2;
    // a request that could never be admitted (its own pages exceed the whole pool) is refused now, not queued forever
    if (s->n_pages > 0 && pages_for(s, n + std::max(1, std::min(max_new, s->max_seq_len - n))) > s->n_pages) return -1;
    Req r;
    r.id = req_id;
    r.prompt.assign(tokens, tokens + n);
    r.max_new = std::max(1, std::min(max_new, s->max_seq_len - n));
    s->reqs.emplace(req_id, std::move(r));
    s->waiting.push_back(req_id);
    s->cv.notify_all();
    return 0;
}

int dl_sched_cancel(void* p, long long req_id) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    auto it = s->reqs.find(req_id);
    if (it == s->reqs.end()) return -1;
    it->second.cancelled = true;
    s->cv.notify_all();
    return 0;
}

// block until there is something to run (1) or the timeout expires (0)
int dl_sched_wait(void* p, int timeout_ms) {
    auto* s = (Sched*)p;
    std::unique_lock<std::mutex> g(s->mu);
    auto has_work = [&] { return !s->waiting.empty() || !s->running.empty(); };
    if (has_work()) return 1;
    s->cv.wait_for(g, std::chrono::milliseconds(timeout_ms), has_work);
    return has_work() ? 1 : 0;
}

static void finish_locked(Sched* s, Req& r) {
    if (r.slot >= 0) {
        s->released.push_back(r.slot);
        s->free_slots.push_back(r.slot);
        r.slot = -1;
    }
    s->pages_reserved -= r.pages;
    r.pages = 0;
    r.state = DONE;
    s->running.erase(std::remove(s->running.begin(), s->running.end(), r.id), s->running.end());
}

// Builds the next batch.  Returns the number of rows (0 = idle).  Rows [*logits_from, n) need
// logits and will each produce one sampled token at commit, in this order.
int dl_sched_next_batch(void* p, long long* req_ids, int* tokens, int* pos, int* slots, int cap, int* logits_from) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    const int R = std::min(cap, s->max_rows);
    // drop cancelled requests
    for (auto it = s->reqs.begin(); it != s->reqs.end(); ++it) {
        Req& r = it->second;
        if (r.cancelled && r.state != DONE) {
            if (r.state == WAITING) s->waiting.erase(std::remove(s->waiting.begin(), s->waiting.end(), r.id), s->waiting.end());
            finish_locked(s, r);
        }
    }
    // admit waiting requests while slots are free
    while (!s->waiting.empty() && !s->free_slots.empty()) {
        Req& r = s->reqs[s->waiting.front()];
        const int need = s->n_pages > 0 ? pages_for(s, (int)r.prompt.size() + r.max_new) : 0;
        if (s->n_pages > 0 && s->pages_reserved + need > s->n_pages) break;    // FIFO: wait until enough pages come back
        r.pages = need;
        s->pages_reserved += need;
        s->waiting.pop_front();
        r.slot = s->free_slots.back();
        s->free_slots.pop_back();
        r.state = PREFILL;
        s->running.push_back(r.id);
    }
    // decode rows (round-robin window if there are more decoders than rows)
    std::vector<Req*> dec, pres;
    int prompt_left = 0;
    for (long long id : s->running) {
        Req& r = s->reqs[id];
        if (r.state == DECODE) dec.push_back(&r);
        else if (r.state == PREFILL) {
            pres.push_back(&r);
            prompt_left += (int)r.prompt.size() - r.n_prefilled;
        }
    }
    int n_dec = (int)dec.size();
    // pending prompts always get a share of the rows (else a full house of decoders starves them)
    const int reserve = pres.empty() ? 0 : std::min(prompt_left, std::max(1, R / 4));
    const int dec_rows = std::max(0, R - reserve);
    if (n_dec > dec_rows) {
        // more decoders than rows: serve a rotating window so everybody advances
        std::rotate(dec.begin(), dec.begin() + (s->rr % dec.size()), dec.end());
        s->rr += dec_rows > 0 ? dec_rows : 1;
        n_dec = dec_rows;
    }
    int n = 0;
    s->batch_sampling.clear();
    // Prompt rows of as many pending requests as fit (admission order, at most prefill_chunk rows per
    // request and step).  Rows that need no logits go first; the LAST row of every prompt that completes
    // in this step is moved to the tail, in front of the decode rows, so that the rows needing logits stay
    // one contiguous range [logits_from, n).  (Rows of a sequence only have to be in increasing position
    // order, not adjacent: the KV rows of the whole batch are written before attention runs.)
    struct TailRow { Req* r; int token, pos; };
    std::vector<TailRow> tail;
    int budget = R - n_dec;
    for (Req* pre : pres) {
        if (budget <= 0) break;
        const int left = (int)pre->prompt.size() - pre->n_prefilled;
        const int take = std::min({left, s->prefill_chunk, budget});
        const bool completes = take == left;
        const int head = completes ? take - 1 : take;
        for (int i = 0; i < head; i++) {
            req_ids[n] = pre->id;
            tokens[n] = pre->prompt[pre->n_prefilled + i];
            pos[n] = pre->pos + i;
            slots[n] = pre->slot;
            n++;
        }
        if (completes) tail.push_back({pre, pre->prompt[pre->n_prefilled + take - 1], pre->pos + take - 1});
        pre->n_prefilled += take;
        pre->pos += take;
        budget -= take;
    }
    *logits_from = n;
    for (const TailRow& t : tail) {                     // last prompt row samples the first new token
        req_ids[n] = t.r->id;
        tokens[n] = t.token;
        pos[n] = t.pos;
        slots[n] = t.r->slot;
        n++;
        s->batch_sampling.push_back(t.r->id);
    }
    for (int i = 0; i < n_dec; i++) {
        Req* r = dec[i];
        req_ids[n] = r->id;
        tokens[n] = r->last_token;
        pos[n] = r->pos;
        slots[n] = r->slot;
        r->pos += 1;
        n++;
        s->batch_sampling.push_back(r->id);
    }
    if (n == 0) *logits_from = 0;
    return n;
}

// number of sampling rows of the batch last returned by next_batch, and their request ids
int dl_sched_sampling_rows(void* p, long long* req_ids, int cap) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    int n = std::min<int>(cap, (int)s->batch_sampling.size());
    for (int i = 0; i < n; i++) req_ids[i] = s->batch_sampling[i];
    return (int)s->batch_sampling.size();
}

// sampled[i] is the token drawn for batch_sampling[i]; stop[i] != 0 ends the request (EOS / stop
// string, decided by the caller).  finished_out[i] = 1 if the request is now done (stop or budget).
int dl_sched_commit(void* p, int n, const int* sampled, const int* stop, int* finished_out) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    if (n != (int)s->batch_sampling.size()) return -1;
    for (int i = 0; i < n; i++) {
        Req& r = s->reqs[s->batch_sampling[i]];
        if (r.state == DONE) { finished_out[i] = 1; continue; }
        r.state = DECODE;
        r.last_token = sampled[i];
        r.n_generated += 1;
        bool fin = stop[i] != 0 || r.n_generated >= r.max_new || r.pos + 1 >= s->max_seq_len;
        finished_out[i] = fin ? 1 : 0;
        if (fin) finish_locked(s, r);
    }
    s->batch_sampling.clear();
    return 0;
}

int dl_sched_pop_released(void* p, int* slots, int cap) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    int n = std::min<int>(cap, (int)s->released.size());
    for (int i = 0; i < n; i++) slots[i] = s->released[i];
    s->released.erase(s->released.begin(), s->released.begin() + n);
    return n;
}

// forget a finished request (frees its prompt memory)
int dl_sched_forget(void* p, long long req_id) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    auto it = s->reqs.find(req_id);
    if (it == s->reqs.end() || it->second.state != DONE) return -1;
    s->reqs.erase(it);
    return 0;
}

// -1 = unknown id, else 0 waiting / 1 prefill / 2 decode / 3 done (finished or cancelled, not yet forgotten)
int dl_sched_state(void* p, long long req_id) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    auto it = s->reqs.find(req_id);
    return it == s->reqs.end() ? -1 : it->second.state;
}

void dl_sched_stats(void* p, int* waiting, int* running, int* free_slots) {
    auto* s = (Sched*)p;
    std::lock_guard<std::mutex> g(s->mu);
    *waiting = (int)s->waiting.size();
    *running = (int)s->running.size();
    *free_slots = (int)s->free_slots.size();
}

}  // extern "C"
