// Multi-threaded writer of random-init `.m` model files (layout: docs/FORMATS.md; reference: converter/writer.py:109-145
// header, src/llm.cpp:447-483 tensor order).  There is no network on the benchmark boxes, so every model is random-init
// weights of the named architecture; the NumPy generator (formats/mfile.py) manages ~0.2 GB/s, which makes the 238 GB
// Llama-3.1-405B file a 20-minute job.  This one fills the file in parallel with pwrite() from a counter-based RNG
// (any chunk can be produced independently), at memory / disk speed.
//
// Statistically the same recipe as the NumPy fast path: Q40 blocks drawn directly in the quantised domain (uniform
// nibbles, scale = +-[0.75, 1.25) * 3*std/8 as fp16), embedding ~ N(0, 1), norm weights 1 + 0.1 * N(0, 1).
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

inline uint64_t splitmix(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

inline float uniform01(uint64_t& s) { return (float)(splitmix(s) >> 40) * (1.0f / 16777216.0f); }

inline float normal(uint64_t& s) {                 // Box-Muller, one value per call
    const float u1 = std::fmax(uniform01(s), 1e-7f), u2 = uniform01(s);
    return sqrtf(-2.0f * logf(u1)) * cosf(6.2831853f * u2);
}

uint16_t f32_to_f16_pos(float f) {                  // positive normal range only (block scales)
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint32_t exp = (x >> 23) & 0xff, man = x & 0x7fffff;
    if (exp < 113) return 0;
    uint32_t h = ((exp - 112) << 10) | (man >> 13);
    const uint32_t rem = man & 0x1fff;
    if (rem > 0x1000 || (rem == 0x1000 && (h & 1))) h++;
    return (uint16_t)h;
}

enum Kind { K_F32_NORMAL, K_F32_NORM_WEIGHT, K_Q40 };
struct Span { int64_t off, bytes; int kind; float scale; };

constexpr int64_t kChunk = 18 * 4 * 65536;          // 4.5 MiB, a multiple of 18 and of 4: chunks hold whole blocks / floats

void fill(uint8_t* buf, int64_t n, const Span& sp, uint64_t seed) {
    uint64_t s = seed;
    if (sp.kind == K_Q40) {
        const int64_t nb = n / 18;
        for (int64_t b = 0; b < nb; b++) {
            uint8_t* o = buf + b * 18;
            const uint64_t r = splitmix(s);
            const float d = ((float)(r >> 40) * (1.0f / 16777216.0f) * 0.5f + 0.75f) * sp.scale;
            uint16_t h = f32_to_f16_pos(d) | ((r & 1) ? 0x8000 : 0);
            memcpy(o, &h, 2);
            const uint64_t a = splitmix(s), c = splitmix(s);
            memcpy(o + 2, &a, 8);
            memcpy(o + 10, &c, 8);
        }
    } else {
        float* f = (float*)buf;
        const int64_t nf = n / 4;
        if (sp.kind == K_F32_NORMAL) for (int64_t i = 0; i < nf; i++) f[i] = normal(s) * sp.scale;
        else for (int64_t i = 0; i < nf; i++) f[i] = 1.0f + 0.1f * normal(s);
    }
}

}  // namespace

extern "C" {

// header_kv: the (key, value) int32 pairs of the header in file order (must contain dim 2, hidden_dim 3, n_layers 4,
// n_heads 5, n_kv_heads 6, vocab_size 9, weights_float_type 13 = 2 (Q40)).  Returns the file size, < 0 on error.
long long dl_write_random_model_q40(const char* path, const int* header_kv, int n_kv, unsigned long long seed, float std_dev,
                                    int n_threads) {
    int64_t dim = 0, hid = 0, layers = 0, heads = 0, kvh = 0, vocab = 0, wt = -1;
    for (int i = 0; i < n_kv; i++) {
        const int k = header_kv[2 * i], v = header_kv[2 * i + 1];
        if (k == 2) dim = v; else if (k == 3) hid = v; else if (k == 4) layers = v; else if (k == 5) heads = v;
        else if (k == 6) kvh = v; else if (k == 9) vocab = v; else if (k == 13) wt = v;
    }
    if (wt != 2 || dim <= 0 || hid <= 0 || layers <= 0 || heads <= 0 || kvh <= 0 || vocab <= 0 || dim % 32 || hid % 32) return -1;
    const int64_t kv_dim = dim * kvh / heads;
    const int64_t header_bytes = 8 + 8 * (int64_t)n_kv;
    auto q40 = [](int64_t rows, int64_t cols) { return rows * (cols / 32) * 18; };
    std::vector<Span> spans;
    int64_t off = header_bytes;
    auto add = [&](int64_t bytes, int kind, float scale) { spans.push_back({off, bytes, kind, scale}); off += bytes; };
    const float qs = std_dev * 3.0f / 8.0f;
    add(vocab * dim * 4, K_F32_NORMAL, 1.0f);
    for (int64_t l = 0; l < layers; l++) {
        add(q40(dim, dim), K_Q40, qs);      // wq
        add(q40(kv_dim, dim), K_Q40, qs);   // wk
        add(q40(kv_dim, dim), K_Q40, qs);   // wv
        add(q40(dim, dim), K_Q40, qs);      // wo
        add(q40(hid, dim), K_Q40, qs);      // w1
        add(q40(dim, hid), K_Q40, qs);      // w2
        add(q40(hid, dim), K_Q40, qs);      // w3
        add(dim * 4, K_F32_NORM_WEIGHT, 0.f);
        add(dim * 4, K_F32_NORM_WEIGHT, 0.f);
    }
    add(dim * 4, K_F32_NORM_WEIGHT, 0.f);
    add(q40(vocab, dim), K_Q40, qs);        // wcls
    const int64_t total = off;

    const std::string tmp = std::string(path) + ".tmp" + std::to_string((long long)getpid());
    const int fd = open(tmp.c_str(), O_CREAT | O_TRUNC | O_WRONLY, 0644);
    if (fd < 0) return -2;
    if (ftruncate(fd, total) != 0) { close(fd); unlink(tmp.c_str()); return -3; }
    std::vector<int32_t> head(2 + 2 * (size_t)n_kv);
    head[0] = 0xA00ABCD;
    head[1] = (int32_t)header_bytes;
    memcpy(head.data() + 2, header_kv, (size_t)n_kv * 8);
    if (pwrite(fd, head.data(), (size_t)header_bytes, 0) != header_bytes) { close(fd); unlink(tmp.c_str()); return -4; }

    // work items: (span, chunk) pairs, claimed with an atomic counter
    struct Item { int span; int64_t off, bytes; };
    std::vector<Item> items;
    for (size_t i = 0; i < spans.size(); i++)
        for (int64_t o = 0; o < spans[i].bytes; o += kChunk) items.push_back({(int)i, o, std::min(kChunk, spans[i].bytes - o)});
    std::atomic<size_t> next{0};
    std::atomic<int> failed{0};
    auto worker = [&] {
        std::vector<uint8_t> buf((size_t)kChunk);
        while (!failed.load()) {
            const size_t i = next.fetch_add(1);
            if (i >= items.size()) break;
            const Item& it = items[i];
            fill(buf.data(), it.bytes, spans[it.span], seed * 0x9E3779B97F4A7C15ULL + (uint64_t)i * 0xD1B54A32D192ED03ULL + 1);
            int64_t done = 0;
            while (done < it.bytes) {
                const ssize_t w = pwrite(fd, buf.data() + done, (size_t)(it.bytes - done), spans[it.span].off + it.off + done);
                if (w <= 0) { failed.store(1); break; }
                done += w;
            }
        }
    };
    const int nt = n_threads > 0 ? n_threads : (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<std::thread> th;
    for (int t = 1; t < nt; t++) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    close(fd);
    if (failed.load()) { unlink(tmp.c_str()); return -5; }
    if (rename(tmp.c_str(), path) != 0) { unlink(tmp.c_str()); return -6; }
    return total;
}

}  // extern "C"
