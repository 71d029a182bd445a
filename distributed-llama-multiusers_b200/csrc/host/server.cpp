// Native front end of `dllama-api`: an epoll HTTP/1.1 server that owns everything around the forward
// pass -- request parsing, chat templating, tokenisation, admission into the continuous-batching
// scheduler, per-request sampler state, streaming detokenisation, stop strings, OpenAI JSON / SSE
// formatting and the socket writes.  The engine thread (Python, holding the GPU engine) only calls
//
//     dl_srv_wait -> dl_srv_next_batch -> [forward + sampling on the device] -> dl_srv_commit
//
// so no Python runs per generated token and no interpreter lock is shared with connection handlers
// (with the Python front end 32 streaming users cost ~5 ms of host time per 3.2 ms GPU step).
//
// Surface parity with the reference (src/dllama-api.cpp:42-170 request parser, :181-191 CORS,
// :222-237 router, :250-313 completions, :320-329 models; src/api-types.hpp DTOs) -- without its
// defects D1-D6: connections are served concurrently, all in-flight requests decode in one batch,
// the chat template is applied, replies are OpenAI `chat.completion` objects or SSE streams of
// `chat.completion.chunk`, text is reassembled to valid UTF-8 before it is put into JSON.
//
// Threads: ONE I/O thread (accept, read, parse, tokenise, submit, write) + the caller's engine thread.
// Shared state (connections' outgoing buffers, generation records) is guarded by Server::mu; the engine
// thread never touches a socket: it appends to a connection's buffer and kicks the I/O thread through
// an eventfd.  Lock order: Server::mu, then the scheduler's own mutex.
#include <arpa/inet.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/epoll.h>
#include <sys/eventfd.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdarg>
#include <ctime>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "json.hpp"

// ---- the rest of the host runtime (text.cpp, scheduler.cpp) ----------------------------------------
extern "C" {
int dl_tok_encode(void* t, const char* text, int text_len, int add_bos, int add_special, int* out, int cap);
int dl_tok_n_eos(void* t);
int dl_tok_eos(void* t, int i);
const char* dl_tok_piece(void* t, int id, int* len);
void* dl_dec_new(void* tok);
void dl_dec_free(void* d);
int dl_dec_push(void* d, int token, char* out, int cap);
void* dl_sampler_new(int vocab, float temperature, float topp, unsigned long long seed);
void dl_sampler_free(void* s);
int dl_sampler_sample(void* s, float* logits);
float dl_sampler_coin(void* s);
int dl_chat_render(int type, const char* eos, int n, const char** roles, const char** msgs, int add_generation_prompt,
                   char* out, int cap, int* public_prompt_len);
void* dl_stop_new(int n, const int* tokens, const char** pieces, int pad_left, int pad_right);
void dl_stop_free(void* d);
void dl_stop_reset(void* d);
int dl_stop_append(void* d, int token, const char* piece, int piece_len);
const char* dl_stop_delta(void* d, int* len);
void* dl_sched_new(int max_rows, int n_slots, int prefill_chunk, int max_seq_len);
void dl_sched_free(void* p);
void dl_sched_set_pages(void* p, int n_pages, int page_size);
int dl_sched_submit(void* p, long long req_id, const int* tokens, int n, int max_new);
int dl_sched_cancel(void* p, long long req_id);
int dl_sched_wait(void* p, int timeout_ms);
int dl_sched_next_batch(void* p, long long* req_ids, int* tokens, int* pos, int* slots, int cap, int* logits_from);
int dl_sched_sampling_rows(void* p, long long* req_ids, int cap);
int dl_sched_commit(void* p, int n, const int* sampled, const int* stop, int* finished_out);
int dl_sched_pop_released(void* p, int* slots, int cap);
int dl_sched_forget(void* p, long long req_id);
int dl_sched_state(void* p, long long req_id);
void dl_sched_stats(void* p, int* waiting, int* running, int* free_slots);
}

namespace {

using dljson::Value;

enum { EOS_MAYBE = 0, EOS_YES = 1, EOS_NO = 2 };
constexpr uint64_t kListenId = 0, kEventId = 1, kFirstConnId = 16;
constexpr size_t kMaxHeader = 64 * 1024, kMaxBody = 32u * 1024 * 1024;
constexpr double kIdleSeconds = 30.0;       // time a client has to deliver its request
constexpr size_t kMaxConnections = 8192;

struct Conn {
    int fd = -1;
    uint64_t id = 0;
    // I/O thread only
    std::string in;
    std::string wbuf;
    size_t woff = 0;
    bool dispatched = false;        // a request of this connection has been routed
    bool sent_continue = false;
    bool want_out = false;          // EPOLLOUT armed
    bool keep_alive = false;        // HTTP/1.1 client that did not ask for "Connection: close"
    double t_open = 0.0;            // connections that never send a complete request are dropped after kIdleSeconds
    // shared (Server::mu)
    std::string out;
    bool close_after = false;       // close once `out` is flushed
    bool done_after = false;        // response complete once `out` is flushed: a keep-alive connection reads the next request
    bool keep = false;              // copy of keep_alive for the thread that finishes the response
    long long req = 0;              // generation in flight for this connection (0 = none)
};

struct Gen {
    long long id = 0;
    uint64_t conn = 0;
    bool stream = false, include_usage = false;
    std::string cid;
    void* sampler = nullptr;
    void* dec = nullptr;
    void* stops = nullptr;
    float temperature = 0.f, topp = 0.f;
    int n_prompt = 0, n_gen = 0;
    std::string text;               // non-streaming answers accumulate here
    const char* finish = nullptr;
    std::string error;
    double t_submit = 0.0;
    bool first_seen = false;
};

struct Server {
    void* tok = nullptr;
    void* sched = nullptr;
    std::string model;
    int chat_type = 0;
    std::string chat_eos;
    int vocab = 0;
    float def_temp = 0.8f, def_topp = 0.9f;
    unsigned long long def_seed = 0;
    int max_default_tokens = 128;
    int max_waiting = 4096;             // admission queue bound (DLLAMA_MAX_WAITING): beyond it requests get 503
    int verbose = 0;
    std::vector<int> eos_ids;
    std::vector<std::string> eos_pieces;
    int max_stop = 1;

    int listen_fd = -1, ep = -1, evfd = -1, port = 0;
    std::thread io;
    std::atomic<bool> stop{false};
    std::atomic<uint64_t> next_conn{kFirstConnId};
    std::atomic<long long> next_req{1};

    std::mutex mu;
    std::unordered_map<uint64_t, Conn> conns;
    std::unordered_map<long long, Gen> gens;
    std::vector<uint64_t> dirty;            // connections with fresh output
    std::vector<long long> cancelled;       // requests whose scheduler entry still has to be forgotten
    // engine thread only
    std::vector<long long> batch_all, batch_ids;
    std::vector<int> scratch_stop, scratch_fin;
    std::vector<char> piece_buf;
    // counters (Server::mu)
    long long n_requests = 0, n_errors = 0, n_steps = 0, n_generated = 0, n_rows = 0, n_engine_errors = 0;
    double ttft_sum = 0.0;
    long long ttft_n = 0;
    std::mt19937_64 rng{std::random_device{}()};
};

double now_s() {
    return std::chrono::duration<double>(std::chrono::system_clock::now().time_since_epoch()).count();
}

void appendf(std::string& o, const char* fmt, ...) {
    char b[512];
    va_list ap;
    va_start(ap, fmt);
    int n = vsnprintf(b, sizeof b, fmt, ap);
    va_end(ap);
    if (n > 0) o.append(b, std::min<size_t>((size_t)n, sizeof b - 1));
}

const char* status_text(int code) {
    switch (code) {
        case 200: return "OK";
        case 204: return "No Content";
        case 400: return "Bad Request";
        case 404: return "Not Found";
        case 413: return "Payload Too Large";
        case 500: return "Internal Server Error";
        case 503: return "Service Unavailable";
    }
    return "OK";
}

// complete response with a body; the connection closes once it is flushed
void http_response(std::string& o, int code, const char* ctype, const std::string& body, bool cors, bool keep = false) {
    appendf(o, "HTTP/1.1 %d %s\r\nServer: dllama-api-b200\r\n", code, status_text(code));
    if (cors) o += "Access-Control-Allow-Origin: *\r\n";
    if (ctype) appendf(o, "Content-Type: %s\r\n", ctype);
    appendf(o, "Content-Length: %zu\r\nConnection: %s\r\n\r\n", body.size(), keep ? "keep-alive" : "close");
    o += body;
}

// marks the response in c.out as complete: the connection closes after the flush, or -- keep-alive client and a
// response with a Content-Length -- goes back to reading requests
void end_response(Conn& c, bool can_keep) {
    if (can_keep && c.keep) c.done_after = true;
    else c.close_after = true;
}

std::string error_json(const std::string& msg, const char* type) {
    std::string b = "{\"error\": {\"message\": ";
    dljson::append_quoted(b, msg);
    b += ", \"type\": ";
    dljson::append_quoted(b, type, strlen(type));
    b += "}}";
    return b;
}

void usage_json(std::string& o, int prompt, int completion) {
    appendf(o, "{\"prompt_tokens\": %d, \"completion_tokens\": %d, \"total_tokens\": %d}", prompt, completion,
            prompt + completion);
}

// one SSE event carrying a chat.completion.chunk (reference DTO: ChatCompletionChunk, src/api-types.hpp:60-84)
void sse_chunk(std::string& o, const Server* s, const Gen& g, const char* delta_key, const char* delta, size_t delta_len,
               const char* role, const char* finish, bool with_usage, const std::string* error) {
    o += "data: {\"id\": ";
    dljson::append_quoted(o, g.cid);
    appendf(o, ", \"object\": \"chat.completion.chunk\", \"created\": %lld, \"model\": ", (long long)time(nullptr));
    dljson::append_quoted(o, s->model);
    o += ", \"choices\": [{\"index\": 0, \"delta\": {";
    if (role) { o += "\"role\": "; dljson::append_quoted(o, role, strlen(role)); }
    if (delta_key) {
        if (role) o += ", ";
        o += "\"";
        o += delta_key;
        o += "\": ";
        dljson::append_quoted(o, delta, delta_len);
    }
    o += "}, \"finish_reason\": ";
    if (finish) dljson::append_quoted(o, finish, strlen(finish)); else o += "null";
    o += "}]";
    if (with_usage) { o += ", \"usage\": "; usage_json(o, g.n_prompt, g.n_gen); }
    if (error) { o += ", \"error\": {\"message\": "; dljson::append_quoted(o, *error); o += ", \"type\": \"server_error\"}"; }
    o += "}\r\n\r\n";
}

void free_gen_handles(Gen& g) {
    if (g.sampler) dl_sampler_free(g.sampler);
    if (g.dec) dl_dec_free(g.dec);
    if (g.stops) dl_stop_free(g.stops);
    g.sampler = g.dec = g.stops = nullptr;
}

void kick(Server* s) {
    uint64_t one = 1;
    ssize_t r = write(s->evfd, &one, sizeof one);
    (void)r;
}

// ---- engine-thread side: text delivery (Server::mu held) -------------------------------------------
void deliver(Server* s, Gen& g, const char* p, int len) {
    if (len <= 0) return;
    if (!g.stream) { g.text.append(p, len); return; }
    auto it = s->conns.find(g.conn);
    if (it == s->conns.end()) return;
    sse_chunk(it->second.out, s, g, "content", p, (size_t)len, nullptr, nullptr, false, nullptr);
    s->dirty.push_back(g.conn);
}

// feed one sampled token through the request's decoder and stop detector; true = generation stops
// (reference: the per-token part of inference_loop, src/app.cpp:369-391, plus chat()'s EosDetector use,
// src/dllama.cpp:182-203)
bool emit(Server* s, Gen& g, int token) {
    if (!g.first_seen) {
        g.first_seen = true;
        s->ttft_sum += now_s() - g.t_submit;
        s->ttft_n++;
    }
    g.n_gen++;
    char* buf = s->piece_buf.data();
    const int n = dl_dec_push(g.dec, token, buf, (int)s->piece_buf.size());
    const int kind = dl_stop_append(g.stops, token, buf, n);
    if (kind == EOS_NO || kind == EOS_YES) {
        int len = 0;
        const char* p = dl_stop_delta(g.stops, &len);
        if (p && len > 0) deliver(s, g, p, len);
        dl_stop_reset(g.stops);
    }
    if (kind == EOS_YES) { g.finish = "stop"; return true; }
    return false;
}

// close a generation: last SSE chunk + [DONE], or the whole chat.completion; frees the record
void finalize(Server* s, long long id, const char* reason) {
    auto git = s->gens.find(id);
    if (git == s->gens.end()) return;
    Gen& g = git->second;
    if (!g.finish) {            // text held back while a stop string might have been forming
        int len = 0;
        const char* p = dl_stop_delta(g.stops, &len);
        if (p && len > 0) deliver(s, g, p, len);
    }
    g.finish = reason;
    const bool failed = strcmp(reason, "error") == 0;
    auto cit = s->conns.find(g.conn);
    if (cit != s->conns.end()) {
        Conn& c = cit->second;
        if (g.stream) {
            sse_chunk(c.out, s, g, nullptr, nullptr, 0, nullptr, reason, g.include_usage, failed ? &g.error : nullptr);
            c.out += "data: [DONE]\r\n\r\n";
        } else if (failed) {
            http_response(c.out, 500, "application/json; charset=utf-8",
                          error_json(g.error.empty() ? "generation failed" : g.error, "server_error"), true, c.keep);
        } else {
            // (reference DTO: ChatCompletion, src/api-types.hpp:86-101; `generated_text` is what the fork's
            // web-ui reads, web-ui/app.js:40)
            std::string b = "{\"id\": ";
            dljson::append_quoted(b, g.cid);
            appendf(b, ", \"object\": \"chat.completion\", \"created\": %lld, \"model\": ", (long long)time(nullptr));
            dljson::append_quoted(b, s->model);
            b += ", \"usage\": ";
            usage_json(b, g.n_prompt, g.n_gen);
            b += ", \"choices\": [{\"index\": 0, \"message\": {\"role\": \"assistant\", \"content\": ";
            dljson::append_quoted(b, g.text);
            b += "}, \"finish_reason\": ";
            dljson::append_quoted(b, reason, strlen(reason));
            b += "}], \"generated_text\": ";
            dljson::append_quoted(b, g.text);
            b += "}";
            http_response(c.out, 200, "application/json; charset=utf-8", b, true, c.keep);
        }
        end_response(c, !g.stream);          // an SSE stream has no length: it ends with the connection
        c.req = 0;
        s->dirty.push_back(c.id);
    }
    if (failed) s->n_errors++;
    if (s->verbose)
        printf("🔶 id=%lld tokens=%d finish=%s\n", id, g.n_gen, reason), fflush(stdout);
    free_gen_handles(g);
    s->gens.erase(git);
}

// ---- I/O thread --------------------------------------------------------------------------------------
void queue_response(Server* s, Conn& c, int code, const char* ctype, const std::string& body, bool cors = true) {
    std::lock_guard<std::mutex> g(s->mu);
    const bool keep = c.keep && code < 400;         // errors close the connection (the request may be half-read)
    http_response(c.out, code, ctype, body, cors, keep);
    if (keep) c.done_after = true; else c.close_after = true;
}

std::string lower(std::string v) {
    for (auto& ch : v) ch = (char)tolower((unsigned char)ch);
    return v;
}

struct HttpRequest {
    std::string method, path;
    size_t content_length = 0;
    bool expect_continue = false;
    bool keep_alive = false;
    size_t header_bytes = 0;
};

// 1 = parsed, 0 = need more bytes, -1 = malformed
int parse_head(const std::string& in, HttpRequest& r) {
    size_t end = in.find("\r\n\r\n");
    if (end == std::string::npos) return in.size() > kMaxHeader ? -1 : 0;
    r.header_bytes = end + 4;
    size_t eol = in.find("\r\n");
    std::string line = in.substr(0, eol);
    size_t a = line.find(' ');
    size_t b = a == std::string::npos ? a : line.find(' ', a + 1);
    if (a == std::string::npos || b == std::string::npos) return -1;
    r.method = line.substr(0, a);
    r.path = line.substr(a + 1, b - a - 1);
    r.keep_alive = line.compare(b + 1, std::string::npos, "HTTP/1.1") == 0;      // 1.1 default; 1.0 closes
    size_t q = r.path.find('?');
    if (q != std::string::npos) r.path.resize(q);
    size_t p = eol + 2;
    while (p < end) {
        size_t e = in.find("\r\n", p);
        if (e == std::string::npos || e > end) e = end;
        size_t colon = in.find(':', p);
        if (colon != std::string::npos && colon < e) {
            std::string key = lower(in.substr(p, colon - p));
            size_t v0 = colon + 1;
            while (v0 < e && (in[v0] == ' ' || in[v0] == '\t')) v0++;
            std::string val = in.substr(v0, e - v0);
            if (key == "content-length") r.content_length = (size_t)strtoull(val.c_str(), nullptr, 10);
            else if (key == "expect" && lower(val).find("100-continue") != std::string::npos) r.expect_continue = true;
            else if (key == "connection") {
                const std::string v = lower(val);
                if (v.find("close") != std::string::npos) r.keep_alive = false;
                else if (v.find("keep-alive") != std::string::npos) r.keep_alive = true;
            }
        }
        p = e + 2;
    }
    return 1;
}

struct Params {
    std::vector<std::string> roles, contents;
    int max_tokens = -1;
    float temperature, topp;
    unsigned long long seed;
    std::vector<std::string> stop;
    bool stream = false, include_usage = false;
};

// (reference: ApiServer::parseRequest + parseChatMessages, src/dllama-api.cpp:291-313, src/api-types.hpp:153-177;
// here every field takes effect)
bool parse_params(const Server* s, const Value& body, Params& p, std::string& err) {
    p.temperature = s->def_temp;
    p.topp = s->def_topp;
    p.seed = s->def_seed;
    const Value* msgs = body.get("messages");
    if (!msgs || msgs->type != Value::ARR) { err = "`messages` must be a list"; return false; }
    for (const Value& m : msgs->arr) {
        const Value* role = m.get("role");
        const Value* content = m.get("content");
        if (m.type != Value::OBJ || !role || !content) { err = "every message needs `role` and `content`"; return false; }
        std::string text;
        if (content->type == Value::STR) text = content->str;
        else if (content->type == Value::ARR) {         // OpenAI content-part arrays: keep the text parts
            for (const Value& part : content->arr) {
                const Value* t = part.get("text");
                if (t && t->type == Value::STR) text += t->str;
            }
        } else if (content->type == Value::NUL) text = "";
        else { err = "message `content` must be a string or a list of parts"; return false; }
        if (role->type != Value::STR) { err = "message `role` must be a string"; return false; }
        p.roles.push_back(role->str);
        p.contents.push_back(std::move(text));
    }
    auto num = [&](const char* k, double& out) {
        const Value* v = body.get(k);
        if (!v || v->is_null()) return false;
        if (v->type != Value::NUM) { err = std::string("`") + k + "` must be a number"; return false; }
        out = v->num;
        return true;
    };
    double d;
    if (const Value* v = body.get("stream")) p.stream = v->truthy();
    if (num("temperature", d)) p.temperature = (float)d;
    if (num("top_p", d)) p.topp = (float)d;
    if (const Value* v = body.get("seed")) {
        if (v->type == Value::NUM) p.seed = (unsigned long long)v->inum;
        else if (!v->is_null()) err = "`seed` must be a number";
    }
    if (num("max_tokens", d)) p.max_tokens = (int)d;
    else if (num("max_completion_tokens", d)) p.max_tokens = (int)d;
    if (!err.empty()) return false;
    if (const Value* so = body.get("stream_options")) {
        const Value* iu = so->get("include_usage");
        if (iu && iu->truthy()) p.include_usage = true;
    }
    if (const Value* st = body.get("stop")) {
        if (st->type == Value::STR) p.stop.push_back(st->str);
        else if (st->type == Value::ARR)
            for (const Value& v : st->arr)
                if (v.type == Value::STR) p.stop.push_back(v.str);
    }
    return true;
}

std::string new_completion_id(Server* s) {
    char b[40];
    uint64_t a, c;
    {
        std::lock_guard<std::mutex> g(s->mu);
        a = s->rng();
        c = s->rng();
    }
    snprintf(b, sizeof b, "cmpl-%016llx%08llx", (unsigned long long)a, (unsigned long long)(c & 0xffffffffu));
    return b;
}

void bad_request(Server* s, Conn& c, const std::string& msg) {
    {
        std::lock_guard<std::mutex> g(s->mu);
        s->n_errors++;
    }
    queue_response(s, c, 400, "application/json; charset=utf-8", error_json(msg, "invalid_request_error"));
}

void handle_completion(Server* s, Conn& c, const char* body, size_t n) {
    Value root;
    std::string err;
    if (n == 0) root.type = Value::OBJ;
    else if (!dljson::Parser(body, n).parse(root, err)) return bad_request(s, c, err);
    if (root.type != Value::OBJ) return bad_request(s, c, "request body must be a JSON object");
    Params p;
    if (!parse_params(s, root, p, err)) return bad_request(s, c, err);
    if (p.roles.empty()) return bad_request(s, c, "Prompt is empty.");
    // chat template -> prompt text -> tokens (reference: ChatTemplateGenerator::generate + Tokenizer::encode)
    const int nm = (int)p.roles.size();
    std::vector<const char*> roles(nm), msgs(nm);
    size_t cap = 256;
    for (int i = 0; i < nm; i++) {
        roles[i] = p.roles[i].c_str();
        msgs[i] = p.contents[i].c_str();
        cap += p.roles[i].size() + p.contents[i].size() + 128 + s->chat_eos.size();
    }
    std::string prompt(cap, '\0');
    int pl = 0;
    int ln = dl_chat_render(s->chat_type, s->chat_eos.c_str(), nm, roles.data(), msgs.data(), 1, &prompt[0], (int)cap, &pl);
    if (ln < 0 || (size_t)ln >= cap) return bad_request(s, c, "chat template rendering failed");
    // NUL bytes inside a message would cut the C strings above short; they never reach the tokenizer
    prompt.resize(ln);
    std::vector<int> tokens(prompt.size() + 2);
    int nt = dl_tok_encode(s->tok, prompt.data(), (int)prompt.size(), 1, 1, tokens.data(), (int)tokens.size());
    if (nt == -2) return bad_request(s, c, "text contains bytes that match no token");
    if (nt <= 0) return bad_request(s, c, "tokenisation failed");
    tokens.resize(nt);

    Gen g;
    g.id = s->next_req++;
    g.conn = c.id;
    g.stream = p.stream;
    g.include_usage = p.include_usage;
    g.cid = new_completion_id(s);
    g.temperature = p.temperature;
    g.topp = p.topp;
    g.n_prompt = nt;
    g.t_submit = now_s();
    const unsigned long long seed = p.seed ? p.seed : (unsigned long long)g.id;
    g.sampler = dl_sampler_new(s->vocab, p.temperature, p.topp, seed);
    g.dec = dl_dec_new(s->tok);
    // stop detector: the tokenizer's EOS tokens (by id and by text) + the request's stop strings
    std::vector<int> stop_ids;
    std::vector<const char*> stop_ptrs;
    int pad = s->max_stop;
    for (size_t i = 0; i < s->eos_ids.size(); i++) {
        if (s->eos_pieces[i].empty()) continue;
        stop_ids.push_back(s->eos_ids[i]);
        stop_ptrs.push_back(s->eos_pieces[i].c_str());
    }
    for (const std::string& st : p.stop) {
        if (st.empty()) continue;
        stop_ids.push_back(-1);
        stop_ptrs.push_back(st.c_str());
        pad = std::max(pad, (int)st.size());
    }
    if (stop_ids.empty()) {                 // a tokenizer without EOS text: match by id only
        for (int id : s->eos_ids) { stop_ids.push_back(id); stop_ptrs.push_back("\x01\x02\x03"); }
    }
    g.stops = dl_stop_new((int)stop_ids.size(), stop_ids.data(), stop_ptrs.data(), pad, pad);
    const int max_tokens = p.max_tokens > 0 ? p.max_tokens : s->max_default_tokens;

    const long long id = g.id;
    int rc;
    {
        int w = 0, run = 0, fr = 0;
        dl_sched_stats(s->sched, &w, &run, &fr);
        if (w >= s->max_waiting) {              // shed load instead of queueing without bound
            free_gen_handles(g);
            {
                std::lock_guard<std::mutex> lk(s->mu);
                s->n_errors++;
            }
            return queue_response(s, c, 503, "application/json; charset=utf-8",
                                  error_json("server is overloaded, try again later", "server_error"));
        }
    }
    {
        std::lock_guard<std::mutex> lk(s->mu);
        rc = dl_sched_submit(s->sched, id, tokens.data(), nt, max_tokens);
        if (rc == 0) {
            s->n_requests++;
            c.req = id;
            if (g.stream) {
                c.out += "HTTP/1.1 200 OK\r\nServer: dllama-api-b200\r\nAccess-Control-Allow-Origin: *\r\n"
                         "Content-Type: text/event-stream; charset=utf-8\r\nCache-Control: no-cache\r\n"
                         "Connection: close\r\n\r\n";
                sse_chunk(c.out, s, g, "content", "", 0, "assistant", nullptr, false, nullptr);
            }
            s->gens.emplace(id, std::move(g));
        }
    }
    if (rc != 0) {
        free_gen_handles(g);
        const char* m = rc == -1 ? "prompt does not fit the context window" : rc == -2 ? "duplicate request id"
                        : rc == -3 ? "empty prompt" : "scheduler error";
        return bad_request(s, c, m);
    }
    if (s->verbose) printf("🔷 POST /v1/chat/completions id=%lld prompt_tokens=%d\n", id, nt), fflush(stdout);
}

void route(Server* s, Conn& c, const HttpRequest& r, const char* body, size_t n) {
    if (r.method == "OPTIONS") {            // CORS preflight (reference: src/dllama-api.cpp:181-191)
        std::lock_guard<std::mutex> g(s->mu);
        c.out += "HTTP/1.1 204 No Content\r\nServer: dllama-api-b200\r\nAccess-Control-Allow-Origin: *\r\n"
                 "Access-Control-Allow-Methods: GET, POST, PUT, DELETE\r\n"
                 "Access-Control-Allow-Headers: Content-Type, Authorization\r\n";
        c.out += c.keep ? "Connection: keep-alive\r\n\r\n" : "Connection: close\r\n\r\n";
        if (c.keep) c.done_after = true; else c.close_after = true;
        return;
    }
    if (r.method == "GET" && r.path == "/v1/models") {
        std::string b = "{\"object\": \"list\", \"data\": [{\"id\": ";
        dljson::append_quoted(b, s->model);
        b += ", \"object\": \"model\", \"created\": 0, \"owned_by\": \"user\"}]}";
        return queue_response(s, c, 200, "application/json; charset=utf-8", b);
    }
    if (r.method == "GET" && (r.path == "/health" || r.path == "/metrics")) {
        int w = 0, run = 0, fr = 0;
        dl_sched_stats(s->sched, &w, &run, &fr);
        std::string b;
        if (r.path == "/health") {
            appendf(b, "{\"status\": \"ok\", \"waiting\": %d, \"running\": %d, \"free_slots\": %d}", w, run, fr);
            return queue_response(s, c, 200, "application/json; charset=utf-8", b);
        }
        long long nr, ne, ns, ng;
        double ttft;
        {
            std::lock_guard<std::mutex> g(s->mu);
            nr = s->n_requests; ne = s->n_errors; ns = s->n_steps; ng = s->n_generated;
            ttft = s->ttft_n ? s->ttft_sum / (double)s->ttft_n : 0.0;
        }
        appendf(b, "# TYPE dllama_requests_total counter\ndllama_requests_total %lld\n", nr);
        appendf(b, "# TYPE dllama_errors_total counter\ndllama_errors_total %lld\n", ne);
        appendf(b, "# TYPE dllama_steps_total counter\ndllama_steps_total %lld\n", ns);
        appendf(b, "# TYPE dllama_generated_tokens_total counter\ndllama_generated_tokens_total %lld\n", ng);
        appendf(b, "# TYPE dllama_running gauge\ndllama_running %d\n# TYPE dllama_waiting gauge\ndllama_waiting %d\n", run, w);
        appendf(b, "# TYPE dllama_ttft_seconds_mean gauge\ndllama_ttft_seconds_mean %.6f\n", ttft);
        return queue_response(s, c, 200, "text/plain; version=0.0.4", b, false);
    }
    if (r.method == "POST" && r.path == "/v1/chat/completions") return handle_completion(s, c, body, n);
    queue_response(s, c, 404, nullptr, "Not Found", false);
}

void close_conn(Server* s, uint64_t id) {
    int fd = -1;
    {
        std::lock_guard<std::mutex> g(s->mu);
        auto it = s->conns.find(id);
        if (it == s->conns.end()) return;
        fd = it->second.fd;
        if (it->second.req) {               // client went away mid-generation: free its slot
            dl_sched_cancel(s->sched, it->second.req);
            s->cancelled.push_back(it->second.req);
        }
        s->conns.erase(it);
    }
    epoll_ctl(s->ep, EPOLL_CTL_DEL, fd, nullptr);
    close(fd);
}

void arm(Server* s, Conn& c, bool out) {
    if (c.want_out == out) return;
    epoll_event ev{};
    ev.events = EPOLLIN | EPOLLRDHUP | (out ? EPOLLOUT : 0);
    ev.data.u64 = c.id;
    epoll_ctl(s->ep, EPOLL_CTL_MOD, c.fd, &ev);
    c.want_out = out;
}

// returns false when the connection was closed
bool flush_conn(Server* s, uint64_t id) {
    Conn* c;
    {
        std::lock_guard<std::mutex> g(s->mu);
        auto it = s->conns.find(id);
        if (it == s->conns.end()) return false;
        c = &it->second;                    // node address is stable; only this thread erases connections
    }
    while (true) {
        bool close_after = false, done_after = false;
        if (c->woff >= c->wbuf.size()) {
            c->wbuf.clear();
            c->woff = 0;
            std::lock_guard<std::mutex> g(s->mu);
            c->wbuf.swap(c->out);
            close_after = c->close_after;
            done_after = c->done_after;
            if (c->wbuf.empty() && done_after) c->done_after = false;
        }
        if (c->wbuf.empty()) {
            if (close_after) { close_conn(s, id); return false; }
            if (done_after) {                       // keep-alive: ready for the next request on this connection
                c->dispatched = false;
                c->sent_continue = false;
                c->t_open = now_s();
            }
            arm(s, *c, false);
            return true;
        }
        ssize_t w = send(c->fd, c->wbuf.data() + c->woff, c->wbuf.size() - c->woff, MSG_NOSIGNAL | MSG_DONTWAIT);
        if (w > 0) { c->woff += (size_t)w; continue; }
        if (w < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) { arm(s, *c, true); return true; }
        if (w < 0 && errno == EINTR) continue;
        close_conn(s, id);
        return false;
    }
}

void on_readable(Server* s, uint64_t id) {
    Conn* c;
    {
        std::lock_guard<std::mutex> g(s->mu);
        auto it = s->conns.find(id);
        if (it == s->conns.end()) return;
        c = &it->second;
    }
    char buf[16384];
    while (true) {
        ssize_t n = recv(c->fd, buf, sizeof buf, MSG_DONTWAIT);
        if (n > 0) {
            if (!c->dispatched) c->in.append(buf, (size_t)n);
            continue;
        }
        if (n == 0) { close_conn(s, id); return; }                        // peer closed
        if (errno == EAGAIN || errno == EWOULDBLOCK) break;
        if (errno == EINTR) continue;
        close_conn(s, id);
        return;
    }
    if (c->dispatched) return;
    HttpRequest r;
    int st = parse_head(c->in, r);
    if (st == 0) return;
    if (st < 0 || r.content_length > kMaxBody) {
        c->dispatched = true;
        queue_response(s, *c, st < 0 ? 400 : 413, "application/json; charset=utf-8",
                       error_json(st < 0 ? "malformed HTTP request" : "request body too large", "invalid_request_error"));
        flush_conn(s, id);
        return;
    }
    if (c->in.size() < r.header_bytes + r.content_length) {
        if (r.expect_continue && !c->sent_continue) {
            c->sent_continue = true;
            {
                std::lock_guard<std::mutex> g(s->mu);
                c->out += "HTTP/1.1 100 Continue\r\n\r\n";
            }
            flush_conn(s, id);
        }
        return;
    }
    c->dispatched = true;
    c->keep_alive = r.keep_alive;
    {
        std::lock_guard<std::mutex> g(s->mu);
        c->keep = r.keep_alive;
    }
    route(s, *c, r, c->in.data() + r.header_bytes, r.content_length);
    std::string().swap(c->in);
    flush_conn(s, id);
}

void io_loop(Server* s) {
    epoll_event evs[128];
    std::vector<uint64_t> todo;
    double last_sweep = now_s();
    while (!s->stop.load(std::memory_order_acquire)) {
        int n = epoll_wait(s->ep, evs, 128, 200);
        for (int i = 0; i < n; i++) {
            const uint64_t id = evs[i].data.u64;
            const uint32_t e = evs[i].events;
            if (id == kListenId) {
                while (true) {
                    int fd = accept4(s->listen_fd, nullptr, nullptr, SOCK_NONBLOCK | SOCK_CLOEXEC);
                    if (fd < 0) break;
                    int one = 1;
                    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
                    const uint64_t cid = s->next_conn++;
                    {
                        std::lock_guard<std::mutex> g(s->mu);
                        if (s->conns.size() >= kMaxConnections) { close(fd); continue; }
                        Conn& c = s->conns[cid];
                        c.fd = fd;
                        c.id = cid;
                        c.t_open = now_s();
                    }
                    epoll_event ev{};
                    ev.events = EPOLLIN | EPOLLRDHUP;
                    ev.data.u64 = cid;
                    epoll_ctl(s->ep, EPOLL_CTL_ADD, fd, &ev);
                }
            } else if (id == kEventId) {
                uint64_t v;
                ssize_t r = read(s->evfd, &v, sizeof v);
                (void)r;
            } else {
                if (e & (EPOLLERR | EPOLLHUP)) { close_conn(s, id); continue; }
                if (e & (EPOLLIN | EPOLLRDHUP)) on_readable(s, id);
                if (e & EPOLLOUT) flush_conn(s, id);
            }
        }
        todo.clear();
        const double now = now_s();
        std::vector<uint64_t> stale;
        {
            std::lock_guard<std::mutex> g(s->mu);
            todo.swap(s->dirty);
            if (now - last_sweep > 1.0) {           // slow or silent clients must not pin connections for ever
                last_sweep = now;
                for (auto& kv : s->conns)
                    if (!kv.second.dispatched && now - kv.second.t_open > kIdleSeconds) stale.push_back(kv.first);
            }
        }
        for (uint64_t id : stale) close_conn(s, id);
        uint64_t last = 0;
        for (uint64_t id : todo) {              // the engine thread queues one entry per chunk
            if (id == last) continue;
            last = id;
            flush_conn(s, id);
        }
    }
}

}  // namespace

// =====================================================================================================
// C ABI (serving/native_server.py)
// =====================================================================================================
extern "C" {

// `tokenizer`: handle from dl_tok_load (must outlive the server).  chat_type: 1 llama2, 2 llama3, 3 deepSeek3.
void* dl_srv_new(void* tokenizer, const char* model_name, int chat_type, const char* chat_eos, int vocab, int max_rows,
                 int n_slots, int prefill_chunk, int max_seq_len, float def_temperature, float def_topp,
                 unsigned long long def_seed, int max_default_tokens, int verbose) {
    auto* s = new Server();
    s->tok = tokenizer;
    s->model = model_name ? model_name : "";
    s->chat_type = chat_type;
    s->chat_eos = chat_eos ? chat_eos : "";
    s->vocab = vocab;
    s->def_temp = def_temperature;
    s->def_topp = def_topp;
    s->def_seed = def_seed;
    s->max_default_tokens = max_default_tokens > 0 ? max_default_tokens : 128;
    s->verbose = verbose;
    s->sched = dl_sched_new(max_rows, n_slots, prefill_chunk, max_seq_len);
    if (const char* mw = getenv("DLLAMA_MAX_WAITING")) s->max_waiting = std::max(0, atoi(mw));
    const int ne = dl_tok_n_eos(tokenizer);
    for (int i = 0; i < ne; i++) {
        int id = dl_tok_eos(tokenizer, i), len = 0;
        const char* p = dl_tok_piece(tokenizer, id, &len);
        s->eos_ids.push_back(id);
        s->eos_pieces.emplace_back(p ? std::string(p, len) : std::string());
        s->max_stop = std::max(s->max_stop, len);
    }
    s->batch_all.resize(max_rows);
    s->batch_ids.resize(max_rows);
    s->scratch_stop.resize(max_rows);
    s->scratch_fin.resize(max_rows);
    s->batch_all.clear();
    s->batch_ids.clear();
    s->piece_buf.resize(65536);
    return s;
}

// binds and starts the I/O thread; returns the bound port (port 0 = ephemeral) or -errno
int dl_srv_listen(void* sv, const char* host, int port) {
    auto* s = (Server*)sv;
    int fd = socket(AF_INET, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    if (fd < 0) return -errno;
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_port = htons((uint16_t)port);
    if (!host || !*host || inet_pton(AF_INET, host, &a.sin_addr) != 1) a.sin_addr.s_addr = htonl(INADDR_ANY);
    if (bind(fd, (sockaddr*)&a, sizeof a) != 0 || listen(fd, 1024) != 0) {
        int e = errno;
        close(fd);
        return -e;
    }
    socklen_t al = sizeof a;
    getsockname(fd, (sockaddr*)&a, &al);
    s->port = ntohs(a.sin_port);
    s->listen_fd = fd;
    s->ep = epoll_create1(EPOLL_CLOEXEC);
    s->evfd = eventfd(0, EFD_NONBLOCK | EFD_CLOEXEC);
    epoll_event ev{};
    ev.events = EPOLLIN;
    ev.data.u64 = kListenId;
    epoll_ctl(s->ep, EPOLL_CTL_ADD, fd, &ev);
    ev.data.u64 = kEventId;
    epoll_ctl(s->ep, EPOLL_CTL_ADD, s->evfd, &ev);
    s->stop = false;
    s->io = std::thread(io_loop, s);
    return s->port;
}

void dl_srv_shutdown(void* sv) {
    auto* s = (Server*)sv;
    if (s->listen_fd < 0) return;
    s->stop.store(true, std::memory_order_release);
    kick(s);
    if (s->io.joinable()) s->io.join();
    {
        std::lock_guard<std::mutex> g(s->mu);
        for (auto& kv : s->conns) close(kv.second.fd);
        s->conns.clear();
    }
    close(s->listen_fd);
    close(s->ep);
    close(s->evfd);
    s->listen_fd = s->ep = s->evfd = -1;
}

// KV page pool of the engine: admission reserves the pages of a request's whole life (see scheduler.cpp)
void dl_srv_set_pages(void* p, int n_pages, int page_size) { dl_sched_set_pages(((Server*)p)->sched, n_pages, page_size); }

void dl_srv_free(void* sv) {
    auto* s = (Server*)sv;
    dl_srv_shutdown(s);
    for (auto& kv : s->gens) free_gen_handles(kv.second);
    dl_sched_free(s->sched);
    delete s;
}

// block until a request is waiting or running (1) or the timeout expires (0)
int dl_srv_wait(void* sv, int timeout_ms) { return dl_sched_wait(((Server*)sv)->sched, timeout_ms); }

// Forms the next batch (reference: the batch-forming half of inference_loop, src/app.cpp:330-367).  Rows
// [*logits_from, n) need logits; they are the *n_sampling rows that dl_srv_commit expects tokens for, in this
// order.  temperature/topp/coin [cap] receive the sampling parameters of those rows (coin: the next draw of the
// request's own xorshift* stream when draw_coins != 0 and temperature != 0).  released: KV slots whose pages
// the engine should free before running the batch.
int dl_srv_next_batch(void* sv, int* tokens, int* pos, int* slots, int cap, int* logits_from, int* n_sampling,
                      float* temperature, float* topp, float* coin, int draw_coins, int* all_greedy, int* released,
                      int released_cap, int* n_released) {
    auto* s = (Server*)sv;
    s->batch_all.resize(cap);
    s->batch_ids.resize(cap);
    const int n = dl_sched_next_batch(s->sched, s->batch_all.data(), tokens, pos, slots, cap, logits_from);
    s->batch_all.resize(n);
    int ns = dl_sched_sampling_rows(s->sched, s->batch_ids.data(), cap);
    ns = std::min(ns, cap);
    s->batch_ids.resize(ns);
    *n_released = dl_sched_pop_released(s->sched, released, released_cap);
    *n_sampling = ns;
    bool greedy = ns > 0;
    std::lock_guard<std::mutex> g(s->mu);
    for (int i = 0; i < ns; i++) {
        auto it = s->gens.find(s->batch_ids[i]);
        if (it == s->gens.end()) {
            temperature[i] = topp[i] = coin[i] = 0.f;
            greedy = false;
            continue;
        }
        Gen& r = it->second;
        temperature[i] = r.temperature;
        topp[i] = r.topp;
        coin[i] = (draw_coins && r.temperature != 0.f) ? dl_sampler_coin(r.sampler) : 0.f;
        if (r.temperature != 0.f) greedy = false;
    }
    *all_greedy = greedy ? 1 : 0;
    // requests cancelled by a disconnect or a failed batch: drop their records once the scheduler let go
    size_t keep = 0;
    for (size_t i = 0; i < s->cancelled.size(); i++) {
        const long long id = s->cancelled[i];
        const int st = dl_sched_state(s->sched, id);
        if (st >= 0 && st != 3 /*DONE*/) { s->cancelled[keep++] = id; continue; }
        if (st == 3) dl_sched_forget(s->sched, id);
        auto it = s->gens.find(id);
        if (it != s->gens.end()) { free_gen_handles(it->second); s->gens.erase(it); }
    }
    s->cancelled.resize(keep);
    return n;
}

// CPU engines / DLLAMA_HOST_SAMPLER=1: sample the rows of the current batch on the host with each request's own
// sampler (reference arithmetic, Sampler::sample, src/tokenizer.cpp:472-502).  logits: [n_sampling][ld], mutated.
int dl_srv_sample_host(void* sv, float* logits, long long ld, int* sampled_out) {
    auto* s = (Server*)sv;
    const int ns = (int)s->batch_ids.size();
    std::vector<void*> samplers(ns, nullptr);
    {
        std::lock_guard<std::mutex> g(s->mu);
        for (int i = 0; i < ns; i++) {
            auto it = s->gens.find(s->batch_ids[i]);
            if (it != s->gens.end()) samplers[i] = it->second.sampler;     // only this thread frees records
        }
    }
    for (int i = 0; i < ns; i++) sampled_out[i] = samplers[i] ? dl_sampler_sample(samplers[i], logits + (long long)i * ld) : 0;
    return ns;
}

// Takes the sampled token of every sampling row of the current batch (n == *n_sampling of dl_srv_next_batch; 0
// after a pure prompt step): detokenises, checks stop strings, streams the text, finishes requests that hit EOS /
// a stop string / their token budget / the context window.  Returns the number of finished requests, < 0 on misuse.
int dl_srv_commit(void* sv, int n, const int* sampled) {
    auto* s = (Server*)sv;
    if (n != (int)s->batch_ids.size()) return -1;
    int n_fin = 0;
    {
        std::lock_guard<std::mutex> g(s->mu);
        s->scratch_stop.assign(n, 0);
        s->scratch_fin.assign(n, 0);
        for (int i = 0; i < n; i++) {
            auto it = s->gens.find(s->batch_ids[i]);
            s->scratch_stop[i] = it == s->gens.end() ? 1 : (emit(s, it->second, sampled[i]) ? 1 : 0);
        }
        if (dl_sched_commit(s->sched, n, sampled, s->scratch_stop.data(), s->scratch_fin.data()) != 0) return -2;
        for (int i = 0; i < n; i++) {
            if (!s->scratch_fin[i]) continue;
            n_fin++;
            const long long id = s->batch_ids[i];
            auto it = s->gens.find(id);
            if (it == s->gens.end()) continue;
            finalize(s, id, it->second.finish ? "stop" : "length");
            dl_sched_forget(s->sched, id);
        }
        s->n_steps++;
        s->n_rows += (long long)s->batch_all.size();
        s->n_generated += n;
    }
    s->batch_ids.clear();
    s->batch_all.clear();
    kick(s);
    return n_fin;
}

// The forward of the current batch failed (KV pool exhausted, CUDA error...): the scheduler has already advanced
// its requests, so they end with finish_reason "error" (HTTP 500 / an `error` object in the last SSE chunk), their
// slots and pages come back, everybody else keeps being served.
int dl_srv_fail_batch(void* sv, const char* message) {
    auto* s = (Server*)sv;
    std::vector<long long> ids;
    for (long long id : s->batch_all)
        if (std::find(ids.begin(), ids.end(), id) == ids.end()) ids.push_back(id);
    const int ns = (int)s->batch_ids.size();
    {
        std::lock_guard<std::mutex> g(s->mu);
        s->n_engine_errors++;
        if (ns > 0) {
            std::vector<int> zeros(ns, 0), ones(ns, 1), fin(ns, 0);
            dl_sched_commit(s->sched, ns, zeros.data(), ones.data(), fin.data());
        }
        for (long long id : ids) {
            dl_sched_cancel(s->sched, id);
            s->cancelled.push_back(id);
            auto it = s->gens.find(id);
            if (it == s->gens.end()) continue;
            it->second.error = message ? message : "generation failed";
            finalize(s, id, "error");
        }
    }
    s->batch_ids.clear();
    s->batch_all.clear();
    kick(s);
    return (int)ids.size();
}

int dl_srv_port(void* sv) { return ((Server*)sv)->port; }

// parse + re-serialise a JSON document with the server's own reader (tests): length written (NUL-terminated),
// -1 = rejected, -2 = `cap` too small
int dl_json_roundtrip(const char* in, int n, char* out, int cap) {
    Value v;
    std::string err, o;
    if (!dljson::Parser(in, (size_t)n).parse(v, err)) return -1;
    dljson::dump(v, o);
    if ((int)o.size() + 1 > cap) return -2;
    memcpy(out, o.data(), o.size());
    out[o.size()] = 0;
    return (int)o.size();
}

// out[0..9]: requests, errors, steps, generated tokens, rows, waiting, running, free slots, open connections,
// failed engine steps
void dl_srv_stats(void* sv, long long* out) {
    auto* s = (Server*)sv;
    int w = 0, r = 0, f = 0;
    dl_sched_stats(s->sched, &w, &r, &f);
    std::lock_guard<std::mutex> g(s->mu);
    out[0] = s->n_requests; out[1] = s->n_errors; out[2] = s->n_steps; out[3] = s->n_generated; out[4] = s->n_rows;
    out[5] = w; out[6] = r; out[7] = f; out[8] = (long long)s->conns.size(); out[9] = s->n_engine_errors;
}

}  // extern "C"
