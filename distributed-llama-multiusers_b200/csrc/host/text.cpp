// Host text runtime: `.t` tokenizer (BPE encode, streaming UTF-8 decode), sampler, chat
// templates, stop-string detector.  C ABI for ctypes (text/native.py).
//
// Behavioural parity targets in the reference (src/tokenizer.cpp): file layout :42-135,
// encode :301-380, streaming decoder :214-299, sampler :25-36,382-502, chat templates :538-612,
// EosDetector :614-699.  The implementations below are independent: the encoder uses a hash map
// and a heap with lazy invalidation (O(n log n)) instead of the reference's O(n^2) rescans, while
// producing the same merges (highest score first, leftmost on ties).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <queue>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

constexpr int kMagicKv = 0x567124;
constexpr int kMagicOld = 0x567123;

enum TokKey { K_VERSION = 0, K_VOCAB = 1, K_MAXLEN = 2, K_BOS = 3, K_EOS = 4, K_PAD = 5, K_CHAT_EOS = 6,
              K_TEMPLATE = 7, K_CHAT_STOP = 8, K_N_EOS = 9 };

struct Tokenizer {
    std::vector<std::string> vocab;
    std::vector<float> scores;
    std::vector<int> eos;
    int bos = -1;
    int max_len = 0;
    bool has_template = false;
    std::string chat_template;
    int n_regular = 0;                                  // ids below bos are "regular", the rest special
    std::unordered_map<std::string, int> regular;       // piece -> id (regular vocab only)
    std::vector<int> special_ids;
};

bool read_exact(FILE* f, void* dst, size_t n) { return fread(dst, 1, n, f) == n; }

Tokenizer* load_tokenizer(const char* path, std::string& err) {
    FILE* f = fopen(path, "rb");
    if (!f) { err = "Failed to open tokenizer file"; return nullptr; }
    auto* t = new Tokenizer();
    auto fail = [&](const char* m) { err = m; fclose(f); delete t; return (Tokenizer*)nullptr; };
    int32_t magic = 0;
    if (!read_exact(f, &magic, 4)) return fail("Cannot read tokenizer magic number");
    int vocab_size = 0;
    if (magic == kMagicOld) {
        int32_t h[5];
        if (!read_exact(f, h, sizeof(h))) return fail("Cannot read tokenizer header");
        vocab_size = h[0]; t->max_len = h[1]; t->bos = h[2]; t->eos.push_back(h[3]);
    } else if (magic == kMagicKv) {
        int32_t header_size = 0;
        if (!read_exact(f, &header_size, 4)) return fail("Cannot read tokenizer header size");
        int n_ints = (header_size - 8) / 4;
        if (n_ints < 0 || n_ints > 4096) return fail("Invalid tokenizer header size");
        std::vector<int32_t> kv(n_ints);
        if (n_ints && !read_exact(f, kv.data(), (size_t)n_ints * 4)) return fail("Cannot read header values");
        int version = -1, tmpl_len = -1, n_eos = 0;
        for (int i = 0; i + 1 < n_ints; i += 2) {
            int key = kv[i], v = kv[i + 1];
            switch (key) {
                case K_VERSION: version = v; break;
                case K_VOCAB: vocab_size = v; break;
                case K_MAXLEN: t->max_len = v; break;
                case K_BOS: t->bos = v; break;
                case K_EOS: case K_CHAT_EOS: t->eos.push_back(v); break;      // backward compatibility
                case K_TEMPLATE: tmpl_len = v; break;
                case K_CHAT_STOP: fseek(f, v, SEEK_CUR); break;                // ignored payload
                case K_PAD: break;
                case K_N_EOS: n_eos = v; break;
                default: return fail("Invalid tokenizer header key");
            }
        }
        if (version != 1) return fail("Old tokenizer version, please regenerate your tokenizer");
        if (tmpl_len > 0) {
            t->chat_template.resize(tmpl_len);
            if (!read_exact(f, &t->chat_template[0], tmpl_len)) return fail("Cannot read chat template from tokenizer file");
            t->has_template = true;
        }
        for (int i = 0; i < n_eos; i++) {
            int32_t id;
            if (!read_exact(f, &id, 4)) return fail("Cannot read eos token id from tokenizer file");
            t->eos.push_back(id);
        }
    } else {
        return fail("Invalid tokenizer file");
    }
    if (t->max_len < 1) return fail("Invalid tokenizer max token length");
    if (vocab_size <= 0) return fail("Invalid tokenizer vocab size");
    t->vocab.resize(vocab_size);
    t->scores.resize(vocab_size);
    for (int i = 0; i < vocab_size; i++) {
        float s; int32_t len;
        if (!read_exact(f, &s, 4) || !read_exact(f, &len, 4) || len < 0) return fail("Cannot read token from tokenizer file");
        t->scores[i] = s;
        t->vocab[i].resize(len);
        if (len && !read_exact(f, &t->vocab[i][0], len)) return fail("Cannot read word from tokenizer file");
    }
    fclose(f);
    // same (fragile) convention as the reference: bos id splits regular / special vocab
    t->n_regular = (t->bos >= 0 && t->bos <= vocab_size) ? t->bos : vocab_size;
    t->regular.reserve((size_t)t->n_regular * 2);
    for (int i = 0; i < t->n_regular; i++) t->regular.emplace(t->vocab[i], i);   // first id wins on duplicates
    for (int i = t->n_regular; i < vocab_size; i++) t->special_ids.push_back(i);
    return t;
}

int find_regular(const Tokenizer* t, const std::string& s) {
    auto it = t->regular.find(s);
    return it == t->regular.end() ? -1 : it->second;
}

int find_special_prefix(const Tokenizer* t, const char* p, size_t remaining) {
    for (int id : t->special_ids) {
        const std::string& s = t->vocab[id];
        if (!s.empty() && s.size() <= remaining && memcmp(s.data(), p, s.size()) == 0) return id;
    }
    return -1;
}

// returns number of tokens, or -1 if `cap` is too small, -2 if trailing bytes matched no token
int encode(const Tokenizer* t, const char* text, size_t len, bool add_bos, bool add_special, int* out, int cap) {
    std::vector<int> toks;
    toks.reserve(len + 1);
    if (add_bos) toks.push_back(t->bos);
    std::string acc;
    for (size_t i = 0; i < len;) {
        if (add_special) {
            int sid = find_special_prefix(t, text + i, len - i);
            if (sid >= 0) { toks.push_back(sid); i += t->vocab[sid].size(); continue; }
        }
        acc.push_back(text[i++]);
        int id = find_regular(t, acc);
        if (id >= 0) { toks.push_back(id); acc.clear(); }
    }
    if (!acc.empty()) return -2;

    // score-ordered pair merging over a doubly linked list + max-heap with lazy invalidation
    const int n = (int)toks.size();
    std::vector<int> prev(n), next(n), stamp(n, 0);
    std::vector<char> alive(n, 1);
    for (int i = 0; i < n; i++) { prev[i] = i - 1; next[i] = (i + 1 < n) ? i + 1 : -1; }
    struct Cand { float score; int left; int id; int ls, rs; };
    auto worse = [](const Cand& a, const Cand& b) {          // heap top = best: highest score, then leftmost
        if (a.score != b.score) return a.score < b.score;
        return a.left > b.left;
    };
    std::priority_queue<Cand, std::vector<Cand>, decltype(worse)> heap(worse);
    std::string buf;
    auto push_pair = [&](int l) {
        int r = next[l];
        if (l < 0 || r < 0) return;
        buf.assign(t->vocab[toks[l]]);
        buf.append(t->vocab[toks[r]]);
        int id = find_regular(t, buf);
        if (id >= 0 && t->scores[id] > -1e10f) heap.push({t->scores[id], l, id, stamp[l], stamp[r]});
    };
    for (int i = 0; i + 1 < n; i++) push_pair(i);
    while (!heap.empty()) {
        Cand c = heap.top();
        heap.pop();
        int l = c.left;
        if (!alive[l]) continue;
        int r = next[l];
        if (r < 0 || stamp[l] != c.ls || stamp[r] != c.rs) continue;     // stale
        toks[l] = c.id;
        stamp[l]++;
        alive[r] = 0;
        next[l] = next[r];
        if (next[r] >= 0) prev[next[r]] = l;
        if (prev[l] >= 0) push_pair(prev[l]);
        push_pair(l);
    }
    int count = 0;
    for (int i = 0; i >= 0 && i < n; i = next[i]) {
        if (count >= cap) return -1;
        out[count++] = toks[i];
    }
    return count;
}

// ---- streaming UTF-8 decoder ------------------------------------------------------------------
struct Decoder {
    const Tokenizer* tok;
    std::string pending;       // bytes of an incomplete trailing UTF-8 sequence
};

int utf8_expected(unsigned char c) {
    if (c <= 0x7f) return 1;
    if (c >= 0xc0 && c <= 0xdf) return 2;
    if (c >= 0xe0 && c <= 0xef) return 3;
    if (c >= 0xf0 && c <= 0xf7) return 4;
    return 0;       // continuation byte or invalid lead
}

// moves every complete code point of `in` to `out`; invalid bytes become U+FFFD; an incomplete but
// so-far-valid tail stays in `in`
void drain_utf8(std::string& in, std::string& out) {
    size_t i = 0;
    const size_t n = in.size();
    while (i < n) {
        unsigned char c = (unsigned char)in[i];
        int need = utf8_expected(c);
        if (need == 0) { out += "\xef\xbf\xbd"; i++; continue; }
        size_t have = 1;
        bool bad = false;
        while (have < (size_t)need && i + have < n) {
            if (((unsigned char)in[i + have] & 0xc0) != 0x80) { bad = true; break; }
            have++;
        }
        if (bad) { out += "\xef\xbf\xbd"; i += have; continue; }    // drop the broken prefix, resync on next byte
        if (have < (size_t)need) break;                              // incomplete tail: keep for the next token
        out.append(in, i, need);
        i += need;
    }
    in.erase(0, i);
}

// ---- sampler ------------------------------------------------------------------------------------
struct Sampler {
    int vocab;
    float temperature, topp;
    uint64_t rng;
    std::vector<std::pair<float, int>> cand;
};

uint32_t rng_u32(uint64_t* s) {          // xorshift* (same generator as the reference, tokenizer.cpp:25-31)
    *s ^= *s >> 12;
    *s ^= *s << 25;
    *s ^= *s >> 27;
    return (uint32_t)((*s * 0x2545F4914F6CDD1DULL) >> 32);
}
float rng_f32(uint64_t* s) { return (rng_u32(s) >> 8) / 16777216.0f; }

void softmax_inplace(float* x, int n) {
    float mx = x[0];
    for (int i = 1; i < n; i++) mx = std::max(mx, x[i]);
    float sum = 0.0f;
    for (int i = 0; i < n; i++) { x[i] = expf(x[i] - mx); sum += x[i]; }
    const float inv = 1.0f / sum;
    for (int i = 0; i < n; i++) x[i] *= inv;
}

int sample(Sampler* s, float* logits) {
    const int n = s->vocab;
    if (s->temperature == 0.0f) {
        int best = 0;
        for (int i = 1; i < n; i++) if (logits[i] > logits[best]) best = i;
        return best;
    }
    for (int i = 0; i < n; i++) logits[i] /= s->temperature;
    softmax_inplace(logits, n);
    const float coin = rng_f32(&s->rng);
    if (s->topp <= 0.0f || s->topp >= 1.0f) {
        float cdf = 0.0f;
        for (int i = 0; i < n; i++) { cdf += logits[i]; if (coin < cdf) return i; }
        return n - 1;
    }
    // nucleus: candidates below (1-p)/(n-1) can never be selected; sort the rest by probability
    const float cutoff = (1.0f - s->topp) / (float)(n - 1);
    s->cand.clear();
    for (int i = 0; i < n; i++) if (logits[i] >= cutoff) s->cand.emplace_back(logits[i], i);
    std::stable_sort(s->cand.begin(), s->cand.end(),
                     [](const std::pair<float, int>& a, const std::pair<float, int>& b) { return a.first > b.first; });
    float cum = 0.0f;
    int last = (int)s->cand.size() - 1;
    for (int i = 0; i < (int)s->cand.size(); i++) {
        cum += s->cand[i].first;
        if (cum > s->topp) { last = i; break; }
    }
    const float r = coin * cum;
    float cdf = 0.0f;
    for (int i = 0; i <= last; i++) { cdf += s->cand[i].first; if (r < cdf) return s->cand[i].second; }
    return s->cand[last].second;
}

// ---- chat templates -----------------------------------------------------------------------------
enum ChatType { CHAT_UNKNOWN = 0, CHAT_LLAMA2 = 1, CHAT_LLAMA3 = 2, CHAT_DEEPSEEK3 = 3 };

int detect_chat(const char* tmpl) {
    if (!tmpl) return CHAT_UNKNOWN;
    if (strstr(tmpl, "[INST]")) return CHAT_LLAMA2;
    if (strstr(tmpl, "<|start_header_id|>")) return CHAT_LLAMA3;
    if (strstr(tmpl, "<\xef\xbd\x9c" "Assistant\xef\xbd\x9c>")) return CHAT_DEEPSEEK3;     // fullwidth bars
    return CHAT_UNKNOWN;
}

std::string render_chat(int type, const std::string& eos, int n, const char** roles, const char** msgs, bool gen,
                        int* public_len) {
    std::string o;
    *public_len = 0;
    auto is = [&](int i, const char* r) { return strcmp(roles[i], r) == 0; };
    if (type == CHAT_LLAMA2) {
        int i = 0;
        if (n >= 2 && is(0, "system") && is(1, "user")) {
            o += std::string("[INST] <<SYS>>\n") + msgs[0] + "\n<</SYS>>\n\n" + msgs[1] + " [/INST]" + eos;
            i = 2;
        }
        for (; i < n; i++) {
            if (is(i, "assistant")) o += std::string(msgs[i]) + eos;
            else if (is(i, "user")) o += std::string("[INST] ") + msgs[i] + " [/INST]" + eos;
        }
    } else if (type == CHAT_LLAMA3) {
        for (int i = 0; i < n; i++)
            o += std::string("<|start_header_id|>") + roles[i] + "<|end_header_id|>\n\n" + msgs[i] + eos;
        if (gen) o += "<|start_header_id|>assistant<|end_header_id|>\n\n";
    } else if (type == CHAT_DEEPSEEK3) {
        const std::string user = "<\xef\xbd\x9c" "User\xef\xbd\x9c>", asst = "<\xef\xbd\x9c" "Assistant\xef\xbd\x9c>";
        int i = 0;
        if (n > 0 && is(0, "system")) { o += msgs[0]; i = 1; }
        for (; i < n; i++) {
            if (is(i, "user")) o += user + msgs[i];
            else if (is(i, "assistant")) o += asst + msgs[i];
        }
        if (gen) { o += asst + "<think>\n"; *public_len = 8; }
    }
    return o;
}

// ---- stop detector --------------------------------------------------------------------------------
enum { EOS_MAYBE = 0, EOS_YES = 1, EOS_NO = 2 };

struct StopDetector {
    std::vector<int> tokens;
    std::vector<std::string> stops;
    int pad_left, pad_right;
    std::string buf;
    int eos_pos = -1;
};

int stop_append(StopDetector* d, int token, const char* piece, int piece_len) {
    if (piece && piece_len > 0) d->buf.append(piece, piece_len);
    for (int tk : d->tokens)
        if (tk == token) { d->eos_pos = (int)d->buf.size(); return EOS_YES; }
    d->eos_pos = -1;
    const int blen = (int)d->buf.size();
    for (const std::string& s : d->stops) {
        const int slen = (int)s.size();
        if (blen > slen + d->pad_left + d->pad_right) continue;
        for (int lo = 0; lo <= d->pad_left; lo++) {
            int n = blen - lo;
            if (n <= 0 || n > slen + d->pad_right) continue;
            if (n > slen) n = slen;
            if (memcmp(d->buf.data() + lo, s.data(), n) == 0) {
                if (n == slen) { d->eos_pos = lo; d->buf.resize(lo); return EOS_YES; }
                return EOS_MAYBE;
            }
        }
    }
    return EOS_NO;
}

}  // namespace

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

void* dl_tok_load(const char* path, char* err, int errlen) {
    std::string e;
    Tokenizer* t = load_tokenizer(path, e);
    if (!t && err && errlen > 0) snprintf(err, errlen, "%s", e.c_str());
    return t;
}
void dl_tok_free(void* t) { delete (Tokenizer*)t; }
int dl_tok_vocab_size(void* t) { return (int)((Tokenizer*)t)->vocab.size(); }
int dl_tok_bos(void* t) { return ((Tokenizer*)t)->bos; }
int dl_tok_n_eos(void* t) { return (int)((Tokenizer*)t)->eos.size(); }
int dl_tok_eos(void* t, int i) { return ((Tokenizer*)t)->eos[i]; }
int dl_tok_n_regular(void* t) { return ((Tokenizer*)t)->n_regular; }
int dl_tok_max_len(void* t) { return ((Tokenizer*)t)->max_len; }
const char* dl_tok_chat_template(void* t) {
    auto* k = (Tokenizer*)t;
    return k->has_template ? k->chat_template.c_str() : nullptr;
}
const char* dl_tok_piece(void* t, int id, int* len) {
    auto* k = (Tokenizer*)t;
    if (id < 0 || id >= (int)k->vocab.size()) { if (len) *len = 0; return nullptr; }
    if (len) *len = (int)k->vocab[id].size();
    return k->vocab[id].data();
}
float dl_tok_score(void* t, int id) { return ((Tokenizer*)t)->scores[id]; }
int dl_tok_is_eos(void* t, int id) {
    for (int e : ((Tokenizer*)t)->eos) if (e == id) return 1;
    return 0;
}
int dl_tok_encode(void* t, const char* text, int text_len, int add_bos, int add_special, int* out, int cap) {
    return encode((Tokenizer*)t, text, (size_t)text_len, add_bos != 0, add_special != 0, out, cap);
}

void* dl_dec_new(void* tok) { auto* d = new Decoder(); d->tok = (Tokenizer*)tok; return d; }
void dl_dec_free(void* d) { delete (Decoder*)d; }
void dl_dec_reset(void* d) { ((Decoder*)d)->pending.clear(); }
// feed one token; writes the newly completed UTF-8 text into out (NUL-terminated); returns its length
int dl_dec_push(void* dv, int token, char* out, int cap) {
    auto* d = (Decoder*)dv;
    const Tokenizer* t = d->tok;
    std::string text;
    if (token == t->bos) {
        // nothing
    } else if (dl_tok_is_eos((void*)t, token)) {
        text.swap(d->pending);          // flush whatever is left, even if incomplete (as the reference does)
    } else if (token >= 0 && token < (int)t->vocab.size()) {
        d->pending += t->vocab[token];
        drain_utf8(d->pending, text);
    }
    int n = (int)std::min<size_t>(text.size(), cap > 0 ? (size_t)cap - 1 : 0);
    if (cap > 0) { memcpy(out, text.data(), n); out[n] = '\0'; }
    return n;
}

void* dl_sampler_new(int vocab, float temperature, float topp, unsigned long long seed) {
    auto* s = new Sampler();
    s->vocab = vocab; s->temperature = temperature; s->topp = topp; s->rng = seed;
    return s;
}
void dl_sampler_free(void* s) { delete (Sampler*)s; }
void dl_sampler_set_temp(void* s, float t) { ((Sampler*)s)->temperature = t; }
void dl_sampler_set_topp(void* s, float p) { ((Sampler*)s)->topp = p; }
void dl_sampler_set_seed(void* s, unsigned long long seed) { ((Sampler*)s)->rng = seed; }
int dl_sampler_sample(void* s, float* logits) { return sample((Sampler*)s, logits); }
// next uniform [0,1) draw of this sampler's stream (for the device sampler: one coin per generated token)
float dl_sampler_coin(void* s) { return rng_f32(&((Sampler*)s)->rng); }
unsigned int dl_rng_u32(unsigned long long* state) { return rng_u32((uint64_t*)state); }
float dl_rng_f32(unsigned long long* state) { return rng_f32((uint64_t*)state); }
void dl_softmax(float* x, int n) { softmax_inplace(x, n); }

int dl_chat_detect(const char* tmpl) { return detect_chat(tmpl); }
// returns the rendered length (may exceed cap-1: call again with a larger buffer)
int dl_chat_render(int type, const char* eos, int n, const char** roles, const char** msgs, int add_generation_prompt,
                   char* out, int cap, int* public_prompt_len) {
    int pl = 0;
    std::string r = render_chat(type, eos ? eos : "", n, roles, msgs, add_generation_prompt != 0, &pl);
    if (public_prompt_len) *public_prompt_len = pl;
    if (cap > 0) {
        size_t c = std::min<size_t>(r.size(), (size_t)cap - 1);
        memcpy(out, r.data(), c);
        out[c] = '\0';
    }
    return (int)r.size();
}

void* dl_stop_new(int n, const int* tokens, const char** pieces, int pad_left, int pad_right) {
    auto* d = new StopDetector();
    for (int i = 0; i < n; i++) { d->tokens.push_back(tokens[i]); d->stops.emplace_back(pieces[i]); }
    d->pad_left = pad_left; d->pad_right = pad_right;
    return d;
}
void dl_stop_free(void* d) { delete (StopDetector*)d; }
void dl_stop_reset(void* d) { ((StopDetector*)d)->buf.clear(); ((StopDetector*)d)->eos_pos = -1; }
int dl_stop_append(void* d, int token, const char* piece, int piece_len) { return stop_append((StopDetector*)d, token, piece, piece_len); }
// text that is safe to show after the last append (NULL/0 when nothing)
const char* dl_stop_delta(void* dv, int* len) {
    auto* d = (StopDetector*)dv;
    if (d->buf.empty() || d->eos_pos == 0) { if (len) *len = 0; return nullptr; }
    if (len) *len = (int)d->buf.size();
    return d->buf.c_str();
}

}  // extern "C"
