// Persistent decode step: ONE cooperative kernel per token (M <= 4 rows) instead of ~160 launches.
//
// Why (measured, profiles/README.md sections 2 and 4): single-stream decode of an 8B model is five
// latency-bound kernels per layer -- each pays launch + dependency release + prologue + pipeline fill
// while its weights stream for 1-5 us -- and at TP=8 two fence/counter/flag hand-offs on top.  Here the
// whole forward pass (the reference's 839 op-steps + 66 socket syncs per token, src/nn/nn-executor.cpp:
// 134-163 over the graph of src/llm.cpp:126-438) is one grid of one CTA per SM:
//
//  * a producer warp walks the weight schedule of ALL phases of ALL layers and keeps a 12-stage ring of
//    bulk async copies (UBLKCP) full; it never waits for a phase boundary, so while the consumers sit in
//    a grid barrier or wait for a peer GPU the next phase's Q40T rows are already landing in shared memory;
//  * 16 consumer warps run the phases: [merge-add + RMSNorm + Q80] -> qkv GEMV (+RoPE, paged KV write) ->
//    split-KV attention -> [split merge + Q80] -> wo GEMV -> [merge-add + RMSNorm + Q80] -> w1|w3 GEMV
//    (SiLU*mul) -> [Q80] -> w2 GEMV -> ... -> final norm -> logits -> greedy pick.  Exact reference
//    arithmetic (int8 dp4a per 32-block, one f32 multiply by d_w*d_x, f32 accumulation);
//  * every CTA owns a contiguous ROW range of each matrix over the full K extent (sub-tile row slices of
//    the Q40T tiles: two bulk copies per k-tile), so no cross-CTA reduction exists anywhere; phases are
//    separated by a device-side grid barrier (one 64-bit atomic + acquire spin per CTA);
//  * tensor parallelism (SYNC_NODE_SLICES + OP_MERGE_ADD, src/nn/nn-network.cpp:537-569, src/llm.cpp:314-322):
//    the wo / w2 epilogue casts each 32-row block of its partial sum to Q80 -- what the reference ships
//    (34 B per 32 values, src/nn/nn-quants.hpp:64-72) -- and stores it as 9 LL words {4 x int8 | fp16
//    scale, epoch} (one aligned 8-byte store each, 72 B per 32 values) into EVERY rank's exchange buffer;
//    the next phase's prologue polls the words themselves, so a hand-off costs one NVLink store latency:
//    no fence, no counter, no flag, and no grid barrier either.  Epochs grow monotonically, a stale word
//    can never match.  All loads/stores to peer memory are plain 8-byte accesses (no TMA on peer memory).
#include <math.h>
#include "epilogue.cuh"
#include "mega_sched.h"

#define MG_NS 12                          // weight ring stages (9216 B each)
#define MG_CONS 512                       // consumer threads (16 warps, 4 groups of 128)
#define MG_THREADS 576                    // + weight producer warp (16) + KV page producer warp (17)
#define MG_MAX_M DL_MEGA_MAX_ROWS
#define MG_ACT_BYTES 98304                // XQ slabs of the current phase | KV page stages + attention merge scratch
#define MG_KV_STAGES 4
#define MG_SMALL_SPLITS 4                 // up to this many splits the wo prologue merges itself
#define MG_PAGE 32
#define MG_LLW 9                          // LL words per 32-value block


// ---- shared memory map (byte offsets from the 1024-aligned base) -------------------------------
#define SM_BARS 0                                         // full[12] empty[12] kv_full[4] kv_empty[4]
#define SM_RING 256
#define SM_ACT (SM_RING + MG_NS * Q40T_TILE_BYTES)        // 110848
#define SM_RED (SM_ACT + MG_ACT_BYTES)                    // [4 groups * kpack][MT][nr]  <= 4*128*4 floats
#define SM_OUT (SM_RED + 4 * 128 * MG_MAX_M * 4)          // out_s [MT][128]
#define SM_QS (SM_OUT + 128 * MG_MAX_M * 4)               // q_s [8][128] f32 (attention)
#define SM_MISC (SM_QS + 8 * 128 * 4)                     // ssq / argmax / w_ml scratch (4 KB)
#define SM_TOTAL (SM_MISC + 4096)

extern void dl_set_error(const char* fmt, ...);

struct MgSpin {
    unsigned long long limit, t0;
    uint32_t* dbg;
    unsigned n;
    DL_DEVINL void init(unsigned long long lim, uint32_t* d) { limit = lim; t0 = 0; n = 0; dbg = d; }
    // called from spin loops; traps (kills the launch with an error instead of hanging the GPU) after `limit` ns
    DL_DEVINL void tick(int code) {
        if (((++n) & 0x3ffu) != 0u || limit == 0ull) return;
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0ull) { t0 = now; return; }
        if (now - t0 > limit) {
            // the launch dies with an error instead of hanging the GPU; `code` says which wait it was
            if (dbg != nullptr) {
                asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(dbg), "r"(0x80000000u | (uint32_t)code * 65536u | blockIdx.x) : "memory");
                __threadfence_system();
            }
            __trap();
        }
    }
    DL_DEVINL void reset() { t0 = 0; n = 0; }
};

DL_DEVINL void mg_mbar_wait(uint64_t* bar, uint32_t parity, MgSpin& sp, int code) {
    while (!mbar_try_wait(bar, parity)) sp.tick(code);
    sp.reset();
}
DL_DEVINL void cons_bar() { asm volatile("bar.sync 1, %0;" ::"n"(MG_CONS) : "memory"); }
DL_DEVINL void cons_kv_bar() { asm volatile("bar.sync 2, %0;" ::"n"(MG_CONS + 32) : "memory"); }
DL_DEVINL void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

DL_DEVINL unsigned long long ld_acquire_gpu_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
DL_DEVINL unsigned long long ld_volatile_u64(const void* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
DL_DEVINL void st_volatile_u64(void* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Grid barrier among the consumer threads of all CTAs (all CTAs are co-resident: cooperative launch).
// The counter only grows: a CTA's ticket t belongs to barrier instance t / G, which completes at (t/G + 1) * G.
DL_DEVINL void grid_sync(unsigned long long* bar, unsigned G, MgSpin& sp) {
    cons_bar();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned long long t = atomicAdd(bar, 1ull);
        const unsigned long long target = (t / G + 1ull) * G;
        while (ld_acquire_gpu_u64(bar) < target) sp.tick(1);
        sp.reset();
    }
    cons_bar();
}

// splits a row with `pos` uses out of the att_splits CTAs the grid provides per (row, kv head)
DL_DEVINL int mg_n_eff(int pos, int att_splits) {
    const int n_pages = pos / MG_PAGE + 1;
    int e;
    if (n_pages <= 32) e = min(MG_SMALL_SPLITS, (n_pages + 1) >> 1);     // <= 1024 tokens: merged by the wo prologue
    else e = (n_pages + 7) >> 3;                                        // long context: >= 8 pages per split
    return max(1, min(e, att_splits));
}

// shared memory is carved with fixed offsets from the dynamic base, so every (non-inlined) phase function finds its
// regions without a pointer table travelling through local memory
extern __shared__ __align__(1024) uint8_t mg_smem[];
struct MgCtx {
    uint64_t *full, *empty, *kv_full, *kv_empty;
    uint8_t *ring, *act;
    float *red, *out_s, *q_s, *misc;
    int tid, cta, G;
    DL_DEVINL MgCtx() {
        full = reinterpret_cast<uint64_t*>(mg_smem + SM_BARS);
        empty = full + MG_NS;
        kv_full = empty + MG_NS;
        kv_empty = kv_full + MG_KV_STAGES;
        ring = mg_smem + SM_RING;
        act = mg_smem + SM_ACT;
        red = reinterpret_cast<float*>(mg_smem + SM_RED);
        out_s = reinterpret_cast<float*>(mg_smem + SM_OUT);
        q_s = reinterpret_cast<float*>(mg_smem + SM_QS);
        misc = reinterpret_cast<float*>(mg_smem + SM_MISC);
        tid = threadIdx.x;
        cta = blockIdx.x;
        G = gridDim.x;
    }
};

// the five matmul phases of a layer (+ logits): shape, row alignment and weights, shared by producer and consumers
enum { PH_QKV = 0, PH_WO = 1, PH_W13 = 2, PH_W2 = 3, PH_CLS = 4 };
struct MgPhase {
    const uint8_t* w;
    int d, n, align;
};
DL_DEVINL MgPhase mg_phase(const DlDecodeArgs& a, const DlDecodeLayer* L, int ph) {
    MgPhase p;
    switch (ph) {
        case PH_QKV: p.w = L->qkv; p.d = a.q_dim + 2 * a.kv_dim; p.n = a.dim; p.align = 2; break;
        case PH_WO: p.w = L->wo; p.d = a.dim; p.n = a.q_dim; p.align = 32; break;
        case PH_W13: p.w = L->w13; p.d = 2 * a.hid; p.n = a.dim; p.align = 2; break;
        case PH_W2: p.w = L->w2; p.d = a.dim; p.n = a.hid; p.align = 32; break;
        default: p.w = a.wcls; p.d = a.vocab_loc; p.n = a.dim; p.align = 2; break;
    }
    return p;
}

// =================================================================================================
// weight producer warp: the schedule of every matmul phase of the whole forward pass
// =================================================================================================
__device__ __noinline__ void produce_matrix(const uint8_t* w, int d, int n, int align, uint32_t& it_io, uint64_t pol,
                                            unsigned long long timeout_ns, uint32_t* dbg) {
    const MgCtx c;
    MgSpin sp;
    sp.init(timeout_ns, dbg);
    const int lane = c.tid & 31;
    const int KT = n >> 7;
    uint32_t it = it_io;
    int R0, R1;
    mg_cta_rows(d, c.G, align, c.cta, R0, R1);
    int r = R0;
    while (r < R1) {
        const int rt = r >> 7, r0 = r & 127;
        const int nr = min(128 - r0, R1 - r);
        const int kpack = mg_kpack(nr);
        const uint8_t* tile0 = w + (size_t)rt * KT * Q40T_TILE_BYTES;
#pragma unroll 1
        for (int kt0 = 0; kt0 < KT; kt0 += kpack) {
            const int nkt = min(kpack, KT - kt0);
            const int s = it % MG_NS;
            if (it >= MG_NS) mg_mbar_wait(&c.empty[s], ((it / MG_NS) & 1) ^ 1, sp, 2);
            uint8_t* st = c.ring + s * Q40T_TILE_BYTES;
            if (lane == 0) mbar_arrive_expect_tx(&c.full[s], (uint32_t)(nkt * nr * 72));
            __syncwarp();
            if (lane < nkt) {
                const uint8_t* t = tile0 + (size_t)(kt0 + lane) * Q40T_TILE_BYTES;
                bulk_g2s_hint(st + lane * nr * 64, t + r0 * 64, (uint32_t)(nr * 64), &c.full[s], pol);
            } else if (lane >= 16 && lane - 16 < nkt) {
                const int j = lane - 16;
                const uint8_t* t = tile0 + (size_t)(kt0 + j) * Q40T_TILE_BYTES;
                bulk_g2s_hint(st + Q40T_NIB_BYTES + j * nr * 8, t + Q40T_SCALE_OFF + r0 * 8, (uint32_t)(nr * 8), &c.full[s], pol);
            }
            it++;
        }
        r += nr;
    }
    it_io = it;
}

DL_DEVINL void weight_producer(const DlDecodeArgs& a) {
    const uint64_t pol = l2_policy_evict_first();
    uint32_t it = 0;
#pragma unroll 1
    for (int l = 0; l <= a.n_layers; l++) {
        const DlDecodeLayer* L = a.layers + (l < a.n_layers ? l : 0);
#pragma unroll 1
        for (int ph = (l < a.n_layers ? PH_QKV : PH_CLS); ph <= (l < a.n_layers ? PH_W2 : PH_CLS); ph++) {
            const MgPhase p = mg_phase(a, L, ph);
            produce_matrix(p.w, p.d, p.n, p.align, it, pol, a.timeout_ns, a.debug);
        }
    }
}

// =================================================================================================
// KV page producer warp: bulk copies of the K and V pages of this CTA's attention work items
// =================================================================================================
__device__ __noinline__ void kv_producer(const DlDecodeArgs& a) {
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int lane = c.tid & 31;
    const int page_bytes = MG_PAGE * a.head_size * 2;
    uint8_t* stages = c.act;
    const int n_items = a.M * a.n_kv_heads * a.att_splits;
    uint32_t ai = 0;
#pragma unroll 1
    for (int l = 0; l < a.n_layers; l++) {
        cons_kv_bar();                 // the consumers passed the grid barrier behind this layer's KV writes
        fence_proxy_async_all();       // generic-proxy stores (this token's K/V) -> async-proxy reads below
        const uint8_t* kc = reinterpret_cast<const uint8_t*>(a.layers[l].k_cache);
        const uint8_t* vc = reinterpret_cast<const uint8_t*>(a.layers[l].v_cache);
        for (int item = c.cta; item < n_items; item += c.G) {
            const int sp_i = item % a.att_splits;
            const int kvh = (item / a.att_splits) % a.n_kv_heads;
            const int m = item / (a.att_splits * a.n_kv_heads);
            const int pos = a.pos[m];
            const int n_eff = mg_n_eff(pos, a.att_splits);
            if (sp_i >= n_eff) continue;
            const int n_pages = pos / MG_PAGE + 1;
            const int pps = (n_pages + n_eff - 1) / n_eff;
            const int pg0 = sp_i * pps, pg1 = min(n_pages, pg0 + pps);
            const int* bt = a.block_table + (size_t)a.slot[m] * a.max_pages;
            for (int i0 = pg0; i0 < pg1; i0 += 32) {
                const int my_page = (i0 + lane < pg1) ? bt[i0 + lane] : 0;
                const int cnt = min(32, pg1 - i0);
                for (int j = 0; j < cnt; j++) {
                    const int page = __shfl_sync(0xffffffffu, my_page, j);
                    const int s = ai % MG_KV_STAGES;
                    if (lane == 0) {
                        if (ai >= MG_KV_STAGES) mg_mbar_wait(&c.kv_empty[s], ((ai / MG_KV_STAGES) & 1) ^ 1, sp, 3);
                        const size_t poff = ((size_t)page * a.n_kv_heads + kvh) * (size_t)page_bytes;
                        mbar_arrive_expect_tx(&c.kv_full[s], 2 * page_bytes);
                        bulk_g2s(stages + (s * 2 + 0) * page_bytes, kc + poff, page_bytes, &c.kv_full[s]);
                        bulk_g2s(stages + (s * 2 + 1) * page_bytes, vc + poff, page_bytes, &c.kv_full[s]);
                    }
                    ai++;
                    __syncwarp();
                }
            }
        }
    }
}

// =================================================================================================
// consumer side
// =================================================================================================
// Q80-quantise 4 consecutive values held by each of the 8 lanes of a 32-value block and store the block into the
// resident activation slab ([kt][M][160] B: 128 int8 | 4 f32 scales | 4 x 8*sum(q)).  All 32 lanes must call.
DL_DEVINL void q80_block8_to_act(float4 v, uint8_t* act, int M, int m, int blk, int sub, bool live) {
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    const float d = amax / 127.0f;
    const float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
    const int q0 = __float2int_rn(v.x * id), q1 = __float2int_rn(v.y * id);
    const int q2 = __float2int_rn(v.z * id), q3 = __float2int_rn(v.w * id);
    int sumq = q0 + q1 + q2 + q3;
    sumq += __shfl_xor_sync(0xffffffffu, sumq, 1);
    sumq += __shfl_xor_sync(0xffffffffu, sumq, 2);
    sumq += __shfl_xor_sync(0xffffffffu, sumq, 4);
    if (!live) return;
    uint8_t* rowp = act + ((size_t)(blk >> 2) * M + m) * XQ_ROW_BYTES;
    const int kb = blk & 3;
    reinterpret_cast<uint32_t*>(rowp + kb * 32)[sub] = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) |
                                                      ((uint32_t)(q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
    if (sub == 0) {
        reinterpret_cast<float*>(rowp + 128)[kb] = __half2float(__float2half_rn(d));
        reinterpret_cast<int*>(rowp + 144)[kb] = 8 * sumq;
    }
}

DL_DEVINL float4 ll_dequant4(unsigned long long wd, unsigned long long ws) {
    const uint32_t p = (uint32_t)wd;
    const float d = __half2float(__ushort_as_half((unsigned short)(ws & 0xffffu)));
    return make_float4((float)(int8_t)(p & 0xff) * d, (float)(int8_t)((p >> 8) & 0xff) * d,
                       (float)(int8_t)((p >> 16) & 0xff) * d, (float)(int8_t)(p >> 24) * d);
}

// ---- prologue 1: x = x_in + sum of the ranks' Q80 partial sums; RMSNorm; Q80 -> act slabs -------------------------
// (OP_MERGE_ADD + OP_INV_RMS + OP_RMS_NORM + OP_CAST, src/llm.cpp:204-239,317-342,394-419)
// Every CTA does the whole row (M*dim <= 16384 floats, L2 reads), so no CTA waits for another one.  The merged row is
// parked in shared memory behind the slabs between the two passes.
// xsrc: residual rows ([M][dim] f32), or NULL = embedding rows of a.tokens (layer 0: OP_EMBEDDING, src/llm.cpp:185-192)
__device__ __noinline__ void prologue_norm(const DlDecodeArgs& a, const float* xsrc, float* x_out, const uint8_t* ll,
                                           uint32_t epoch, const float* norm_w) {
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int M = a.M, dim = a.dim, tid = c.tid;
    const int cpr = dim >> 2;                              // float4 chunks per row
    const int total = M * cpr;
    const int nblk = dim >> 5;
    const int N = (ll != nullptr) ? a.world : 0;
    float4* xs = reinterpret_cast<float4*>(c.act + (size_t)(dim >> 7) * M * XQ_ROW_BYTES);     // [M][dim] f32
    float ss[MG_MAX_M];
#pragma unroll
    for (int m = 0; m < MG_MAX_M; m++) ss[m] = 0.0f;
#pragma unroll 1
    for (int ch = tid; ch < total; ch += MG_CONS) {
        const int m = ch / cpr, e4 = ch - m * cpr;
        // the residual rows are rewritten by other CTAs every phase: read them from L2, never from a stale L1 line
        float4 x4 = (xsrc != nullptr) ? __ldcg(reinterpret_cast<const float4*>(xsrc + (size_t)m * dim) + e4)
                                      : reinterpret_cast<const float4*>(a.embedding + (size_t)a.tokens[m] * dim)[e4];
        if (N > 0) {
            const int blk = e4 >> 3, sub = e4 & 7;
            // word (src, m, blk, w) lives at index ((src*4 + m)*nblk + blk)*9 + w
            const unsigned long long* base = reinterpret_cast<const unsigned long long*>(ll) + ((size_t)m * nblk + blk) * MG_LLW;
            const size_t sstride = (size_t)MG_MAX_M * nblk * MG_LLW;
            unsigned long long wd[DL_MAX_PEERS], ws[DL_MAX_PEERS];
            bool ok;
            do {
                ok = true;
#pragma unroll
                for (int s = 0; s < DL_MAX_PEERS; s++) {
                    if (s < N) {
                        wd[s] = ld_volatile_u64(base + s * sstride + sub);
                        ws[s] = ld_volatile_u64(base + s * sstride + 8);
                    }
                }
#pragma unroll
                for (int s = 0; s < DL_MAX_PEERS; s++)
                    if (s < N) ok = ok && ((uint32_t)(wd[s] >> 32) == epoch) && ((uint32_t)(ws[s] >> 32) == epoch);
                if (!ok) sp.tick(4);
            } while (!ok);
            sp.reset();
#pragma unroll
            for (int s = 0; s < DL_MAX_PEERS; s++) {
                if (s < N) {
                    const float4 p = ll_dequant4(wd[s], ws[s]);
                    x4.x += p.x; x4.y += p.y; x4.z += p.z; x4.w += p.w;
                }
            }
        }
        if (x_out != nullptr) reinterpret_cast<float4*>(x_out + (size_t)m * dim)[e4] = x4;
        xs[ch] = x4;
        const float sq = x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
#pragma unroll
        for (int q = 0; q < MG_MAX_M; q++) ss[q] += (m == q) ? sq : 0.0f;
    }
    float* ssq_w = c.misc;                // [16 warps][4]
    const int warp = tid >> 5, lane = tid & 31;
#pragma unroll
    for (int m = 0; m < MG_MAX_M; m++) {
        const float t = warp_sum(ss[m]);
        if (lane == 0) ssq_w[warp * MG_MAX_M + m] = t;
    }
    cons_bar();
    float inv[MG_MAX_M];
#pragma unroll
    for (int m = 0; m < MG_MAX_M; m++) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < MG_CONS / 32; w++) t += ssq_w[w * MG_MAX_M + m];
        inv[m] = 1.0f / sqrtf(t / (float)dim + a.eps);
    }
#pragma unroll 1
    for (int ch0 = 0; ch0 < total; ch0 += MG_CONS) {       // total is a multiple of 32: warps are all-live or all-dead
        const int ch = ch0 + tid;
        if (ch - lane >= total) break;
        const bool live = ch < total;
        const int m = live ? ch / cpr : 0, e4 = live ? ch - m * cpr : 0;
        float im = inv[0];
#pragma unroll
        for (int q = 1; q < MG_MAX_M; q++) im = (m == q) ? inv[q] : im;
        const float4 w4 = reinterpret_cast<const float4*>(norm_w)[e4];
        const float4 x4 = live ? xs[ch] : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 y;
        y.x = w4.x * (im * x4.x); y.y = w4.y * (im * x4.y);
        y.z = w4.z * (im * x4.z); y.w = w4.w * (im * x4.w);
        q80_block8_to_act(y, c.act, M, m, e4 >> 3, e4 & 7, live);
    }
    cons_bar();
}

// ---- prologue 2: merge the attention splits, Q80 -> act slabs (K = q_dim)  (OP_CAST y->yq_slice, src/llm.cpp:296-301)
__device__ __noinline__ void prologue_att(const DlDecodeArgs& a, bool merged) {
    const MgCtx c;
    const int M = a.M, cpr = a.q_dim >> 2, total = M * cpr, hs = a.head_size, tid = c.tid, lane = tid & 31;
#pragma unroll 1
    for (int ch0 = 0; ch0 < total; ch0 += MG_CONS) {
        const int ch = ch0 + tid;
        if (ch - lane >= total) break;
        const bool live = ch < total;
        const int m = live ? ch / cpr : 0, e4 = live ? ch - m * cpr : 0;
        const int head = (e4 * 4) / hs, dd = (e4 * 4) - head * hs;
        const int n_eff = merged ? 1 : mg_n_eff(a.pos[m], a.att_splits);
        const size_t base = ((size_t)m * a.n_heads + head) * a.max_splits;
        float Mx = -INFINITY;
        for (int s = 0; s < n_eff; s++) Mx = fmaxf(Mx, __ldcg(&a.att_ml[(base + s) * 2]));
        const float Mu = (Mx == -INFINITY) ? 0.0f : Mx;
        float L = 0.0f;
        float4 O = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s = 0; s < n_eff; s++) {
            const float e = __expf(__ldcg(&a.att_ml[(base + s) * 2]) - Mu);
            L += __ldcg(&a.att_ml[(base + s) * 2 + 1]) * e;
            const float4 o = __ldcg(reinterpret_cast<const float4*>(a.att_o + (base + s) * hs + dd));
            O.x += o.x * e; O.y += o.y * e; O.z += o.z * e; O.w += o.w * e;
        }
        const float il = (L > 0.0f) ? 1.0f / L : 0.0f;
        O.x *= il; O.y *= il; O.z *= il; O.w *= il;
        q80_block8_to_act(O, c.act, M, m, e4 >> 3, e4 & 7, live);
    }
    cons_bar();
}

// ---- prologue 3: hidden activation f32 -> Q80 act slabs (K = hid)  (OP_CAST d->dq, src/llm.cpp:367-374)
__device__ __noinline__ void prologue_hidden(const DlDecodeArgs& a) {
    const MgCtx c;
    const int M = a.M, cpr = a.hid >> 2, total = M * cpr, tid = c.tid, lane = tid & 31;
#pragma unroll 1
    for (int ch0 = 0; ch0 < total; ch0 += MG_CONS) {
        const int ch = ch0 + tid;
        if (ch - lane >= total) break;
        const bool live = ch < total;
        const int m = live ? ch / cpr : 0, e4 = live ? ch - m * cpr : 0;
        float4 h4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) h4 = __ldcg(reinterpret_cast<const float4*>(a.h + (size_t)m * a.hid) + e4);
        q80_block8_to_act(h4, c.act, M, m, e4 >> 3, e4 & 7, live);
    }
    cons_bar();
}

// ---- epilogue: RoPE + q store + paged KV write  (OP_ROPE_LLAMA + OP_SHIFT, src/llm.cpp:259-286) -------------------
DL_DEVINL void epi_qkv(const DlDecodeArgs& a, const MgCtx& c, const DlDecodeLayer* L, int row, int nr) {
    const int M = a.M, hs = a.head_size;
#pragma unroll 1
    for (int idx = c.tid; idx < M * (nr >> 1); idx += MG_CONS) {
        const int m = idx / (nr >> 1), pr = idx - m * (nr >> 1);
        const int grow = row + 2 * pr;                         // even row of the pair
        float v0 = c.out_s[m * 128 + 2 * pr], v1 = c.out_s[m * 128 + 2 * pr + 1];
        const int region = (grow < a.q_dim) ? 0 : ((grow < a.q_dim + a.kv_dim) ? 1 : 2);
        const int rrow = (region == 0) ? grow : ((region == 1) ? grow - a.q_dim : grow - a.q_dim - a.kv_dim);
        const int hd = rrow % hs, kvh = rrow / hs;
        const int p = a.pos[m];
        if (region < 2) {
            float sn, cs;
            sincosf((float)p * a.rope_freqs[hd >> 1], &sn, &cs);
            const float r0v = v0 * cs - v1 * sn, r1v = v0 * sn + v1 * cs;
            v0 = r0v; v1 = r1v;
        }
        if (region == 0) {
            *reinterpret_cast<float2*>(a.q + (size_t)m * a.q_dim + rrow) = make_float2(v0, v1);
        } else {
            const int page = a.block_table[(size_t)a.slot[m] * a.max_pages + p / MG_PAGE];
            const size_t off = (((size_t)page * a.n_kv_heads + kvh) * MG_PAGE + (p % MG_PAGE)) * hs + hd;
            __half* base = reinterpret_cast<__half*>(region == 1 ? L->k_cache : L->v_cache);
            *reinterpret_cast<__half2*>(base + off) = __floats2half2_rn(v0, v1);
        }
    }
    fence_proxy_async_all();          // the KV rows just written are read by bulk copies after the next grid barrier
}

// ---- epilogue: Q80 cast of a partial sum (32-row blocks), residual add (1 GPU) or LL push to every rank ----------
// (OP_CAST y->ZQ slice + SYNC_NODE_SLICES + OP_MERGE_ADD, src/llm.cpp:308-322,381-387)
// x_emb: the residual comes from the embedding rows (layer 0 on one GPU) instead of a.x[0]
DL_DEVINL void epi_residual(const DlDecodeArgs& a, const MgCtx& c, int row, int nr, bool x_emb, uint8_t* const* ll, uint32_t epoch) {
    const int M = a.M, warp = c.tid >> 5, lane = c.tid & 31;
    const int nb = nr >> 5;                                    // 32-row blocks in this piece
    const int nblk = a.dim >> 5;
#pragma unroll 1
    for (int item = warp; item < M * nb; item += MG_CONS / 32) {
        const int m = item / nb, b = item - m * nb;
        const int grow = row + b * 32 + lane;
        const float val = c.out_s[m * 128 + b * 32 + lane];
        const float amax = warp_max(fabsf(val));
        const float d = amax / 127.0f;
        const float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
        const int q = __float2int_rn(val * id);
        const __half d16 = __float2half_rn(d);
        if (ll == nullptr) {
            const float vq = a.q80_sync ? (float)q * __half2float(d16) : val;
            const float xi = x_emb ? a.embedding[(size_t)a.tokens[m] * a.dim + grow] : __ldcg(&a.x[0][(size_t)m * a.dim + grow]);
            a.x[0][(size_t)m * a.dim + grow] = xi + vq;
        } else {
            uint32_t bts = ((uint32_t)q & 0xffu) << (8 * (lane & 3));
            bts |= __shfl_xor_sync(0xffffffffu, bts, 1);
            bts |= __shfl_xor_sync(0xffffffffu, bts, 2);
            const uint32_t wj = __shfl_sync(0xffffffffu, bts, (lane & 7) * 4);
            const uint32_t payload = (lane < 8) ? wj : (uint32_t)__half_as_ushort(d16);
            if (lane < MG_LLW) {
                const unsigned long long word = ((unsigned long long)epoch << 32) | payload;
                const size_t widx = ((((size_t)a.rank * MG_MAX_M + m) * nblk) + (grow >> 5)) * MG_LLW + lane;
                for (int p = 0; p < a.world; p++)
                    st_volatile_u64(reinterpret_cast<unsigned long long*>(ll[p]) + widx, word);
            }
        }
    }
}

// ---- epilogue: SiLU(gate) * up -> hidden f32  (OP_SILU + OP_MUL, src/llm.cpp:355-366) --------------------------------
DL_DEVINL void epi_silu(const DlDecodeArgs& a, const MgCtx& c, int row, int nr) {
    const int M = a.M;
#pragma unroll 1
    for (int idx = c.tid; idx < M * (nr >> 1); idx += MG_CONS) {
        const int m = idx / (nr >> 1), pr = idx - m * (nr >> 1);
        const float gt = c.out_s[m * 128 + 2 * pr], up = c.out_s[m * 128 + 2 * pr + 1];
        const float hv = (a.act_gelu ? act_gelu(gt) : act_silu(gt)) * up;
        a.h[(size_t)m * a.hid + ((row >> 1) + pr)] = hv;
    }
}

DL_DEVINL void argmax_combine2(float& best, int& bi, float ov, int oi) {
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
}

// ---- epilogue: logits slice -> every rank, or (greedy) the running best of this CTA in shared memory ----------------
// (logits gather SYNC_NODE_SLICES_EXCEPT_ROOT, src/llm.cpp:426-432; Sampler::sample temperature == 0, src/tokenizer.cpp:472-502)
#define MG_BEST_OFF 512      // floats into misc: [4 warps][4][2] scratch, then [4][2] running best
DL_DEVINL void epi_logits(const DlDecodeArgs& a, const MgCtx& c, int row, int nr) {
    const int M = a.M, tid = c.tid;
    if (!a.greedy) {
#pragma unroll 1
        for (int idx = tid; idx < (M - a.row0) * nr; idx += MG_CONS) {
            const int mm = idx / nr, lr = idx - mm * nr;
            const float v = c.out_s[(mm + a.row0) * 128 + lr];
            const size_t off = (size_t)mm * a.vocab + (size_t)a.rank * a.vocab_loc + row + lr;
            for (int p = 0; p < a.world; p++) a.logits[p][off] = v;
        }
        return;
    }
    float* scr = c.misc + MG_BEST_OFF;
    float* run = scr + 4 * MG_MAX_M * 2;
    const int warp = tid >> 5, lane = tid & 31;
    if (warp < 4) {
        for (int m = 0; m < M; m++) {
            float bv = (tid < nr) ? c.out_s[m * 128 + tid] : -INFINITY;
            int bi = (tid < nr) ? a.rank * a.vocab_loc + row + tid : 0x7fffffff;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                argmax_combine2(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
            if (lane == 0) { scr[(warp * MG_MAX_M + m) * 2] = bv; scr[(warp * MG_MAX_M + m) * 2 + 1] = __int_as_float(bi); }
        }
    }
    cons_bar();
    if (tid < M) {
        float bv = run[tid * 2];
        int bi = __float_as_int(run[tid * 2 + 1]);
        for (int w = 0; w < 4; w++) argmax_combine2(bv, bi, scr[(w * MG_MAX_M + tid) * 2], __float_as_int(scr[(w * MG_MAX_M + tid) * 2 + 1]));
        run[tid * 2] = bv;
        run[tid * 2 + 1] = __int_as_float(bi);
    }
}

// ---- the matmul phase: this CTA's rows of a [d][n] Q40T matrix against the resident act slabs ------------------
// After each piece out_s[m*128 + local row] holds the f32 products and the phase's epilogue runs on it.
template <int MT>
__device__ __noinline__ void gemv_phase(const DlDecodeArgs& a, int ph, int layer, uint32_t& it_io, uint32_t epoch) {
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const DlDecodeLayer* Lw = a.layers + (ph == PH_CLS ? 0 : layer);
    const MgPhase P = mg_phase(a, Lw, ph);
    const int M = a.M;
    const int KT = P.n >> 7;
    const int tid = c.tid, g = tid >> 7, t = tid & 127, lane = tid & 31;
    uint32_t it = it_io;
    int R0, R1;
    mg_cta_rows(P.d, c.G, P.align, c.cta, R0, R1);
    int r = R0;
#pragma unroll 1
    while (r < R1) {
        const int r0 = r & 127;
        const int nr = min(128 - r0, R1 - r);
        const int kpack = mg_kpack(nr);
        const int nch = (KT + kpack - 1) / kpack;
        const int j = t / nr, lr = t - j * nr;
        const bool mine = j < kpack;
        const int rot = (r0 + lr) >> 1;
        float acc[MT];
#pragma unroll
        for (int m = 0; m < MT; m++) acc[m] = 0.0f;
#pragma unroll 1
        for (int ci = 0; ci < nch; ci++) {
            const uint32_t itc = it + ci;
            if ((int)(itc & 3u) != g) continue;
            const int s = itc % MG_NS;
            mg_mbar_wait(&c.full[s], (itc / MG_NS) & 1, sp, 5);
            const int kt = ci * kpack + j;
            if (mine && kt < KT) {
                const uint8_t* st = c.ring + s * Q40T_TILE_BYTES;
                const uint8_t* nibp = st + (j * nr + lr) * 64;
                const uint8_t* scp = st + Q40T_NIB_BYTES + (j * nr + lr) * 8;
                const uint8_t* actp = c.act + (size_t)kt * M * XQ_ROW_BYTES;
#pragma unroll
                for (int kb = 0; kb < 4; kb++) {
                    const uint4 nb = *reinterpret_cast<const uint4*>(nibp + (((kb + rot) & 3) << 4));
                    const float dw = __half2float(*reinterpret_cast<const __half*>(scp + kb * 2));
                    const int lo0 = nb.x & 0x0F0F0F0F, hi0 = (nb.x >> 4) & 0x0F0F0F0F;
                    const int lo1 = nb.y & 0x0F0F0F0F, hi1 = (nb.y >> 4) & 0x0F0F0F0F;
                    const int lo2 = nb.z & 0x0F0F0F0F, hi2 = (nb.z >> 4) & 0x0F0F0F0F;
                    const int lo3 = nb.w & 0x0F0F0F0F, hi3 = (nb.w >> 4) & 0x0F0F0F0F;
#pragma unroll
                    for (int m = 0; m < MT; m++) {
                        if (m < M) {
                            const uint8_t* xr = actp + m * XQ_ROW_BYTES;
                            const int4 xl = *reinterpret_cast<const int4*>(xr + kb * 32);
                            const int4 xh = *reinterpret_cast<const int4*>(xr + kb * 32 + 16);
                            int is0 = __dp4a(lo0, xl.x, 0), is1 = __dp4a(hi0, xh.x, 0);
                            is0 = __dp4a(lo1, xl.y, is0); is1 = __dp4a(hi1, xh.y, is1);
                            is0 = __dp4a(lo2, xl.z, is0); is1 = __dp4a(hi2, xh.z, is1);
                            is0 = __dp4a(lo3, xl.w, is0); is1 = __dp4a(hi3, xh.w, is1);
                            const float dx = reinterpret_cast<const float*>(xr + 128)[kb];
                            const int s8 = reinterpret_cast<const int*>(xr + 144)[kb];
                            acc[m] += (float)(is0 + is1 - s8) * (dw * dx);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&c.empty[s]);
        }
        it += nch;
        if (mine) {
#pragma unroll
            for (int m = 0; m < MT; m++) c.red[((g * kpack + j) * MT + m) * nr + lr] = acc[m];
        }
        cons_bar();
#pragma unroll 1
        for (int idx = tid; idx < MT * nr; idx += MG_CONS) {
            const int m = idx / nr, l2 = idx - m * nr;
            float vs = 0.0f;
            for (int q = 0; q < 4 * kpack; q++) vs += c.red[(q * MT + m) * nr + l2];
            c.out_s[m * 128 + l2] = vs;
        }
        cons_bar();
        switch (ph) {
            case PH_QKV: epi_qkv(a, c, Lw, r, nr); break;
            case PH_WO: epi_residual(a, c, r, nr, layer == 0, a.world > 1 ? a.ll_att : nullptr, epoch); break;
            case PH_W13: epi_silu(a, c, r, nr); break;
            case PH_W2: epi_residual(a, c, r, nr, false, a.world > 1 ? a.ll_ffn : nullptr, epoch); break;
            default: epi_logits(a, c, r, nr); break;
        }
        cons_bar();
        r += nr;
    }
    it_io = it;
}

// =================================================================================================
// attention phase: split-KV flash-decode over the paged fp16 cache (OP_MULTIHEAD_ATT, nn-cpu-ops.cpp:749-784)
// item = (row, kv head, split); NGA groups of 4 warps take the item's pages round-robin, the G query heads of the
// kv head share every K/V row.  Partials (unnormalised O, max, sum) go to att_o / att_ml[.., split].
// =================================================================================================
DL_DEVINL float warp_transpose_reduce32_m(float (&v)[32], int lane) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const bool upper = (lane & o) != 0;
#pragma unroll
        for (int k = 0; k < o; k++) {
            const float send = upper ? v[k] : v[k + o];
            const float keep = upper ? v[k + o] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0];
}

template <int HS, int NCH>
__device__ __noinline__ void attention_phase(const DlDecodeArgs& a, uint32_t& ai_io) {
    constexpr int DPL = HS / 32;
    constexpr int GMAX = NCH * 4;
    constexpr int NGA = (NCH == 1) ? 4 : 2;                 // warp groups working on pages concurrently
    constexpr int AW = NGA * 4;                             // attention warps
    constexpr int PAGE_BYTES = MG_PAGE * HS * 2;
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int tid = c.tid, warp = tid >> 5, lane = tid & 31;
    const int G = a.n_heads / a.n_kv_heads;
    uint8_t* stages = c.act;
    float* w_acc = reinterpret_cast<float*>(c.act + MG_KV_STAGES * 2 * PAGE_BYTES);   // [AW][GMAX][HS]
    float* w_ml = c.misc;                                                             // [AW][GMAX][2]
    float* q_s = c.q_s;
    const int n_items = a.M * a.n_kv_heads * a.att_splits;
    uint32_t ai = ai_io;
#pragma unroll 1
    for (int item = c.cta; item < n_items; item += c.G) {
        const int sp_i = item % a.att_splits;
        const int kvh = (item / a.att_splits) % a.n_kv_heads;
        const int m = item / (a.att_splits * a.n_kv_heads);
        const int pos = a.pos[m];
        const int n_eff = mg_n_eff(pos, a.att_splits);
        if (sp_i >= n_eff) continue;
        const int T = pos + 1;
        const int n_pages = pos / MG_PAGE + 1;
        const int pps = (n_pages + n_eff - 1) / n_eff;
        const int pg0 = sp_i * pps, pg1 = min(n_pages, pg0 + pps);
        const int np = max(0, pg1 - pg0);
        const float scale = rsqrtf((float)HS);
        for (int i = tid; i < GMAX * HS; i += MG_CONS) {
            const int h = i / HS, dd = i - h * HS;
            q_s[i] = (h < G) ? __ldcg(&a.q[(size_t)m * a.q_dim + (size_t)(kvh * G + h) * HS + dd]) * scale : 0.0f;
        }
        cons_bar();
        if (warp < AW) {
            const int ga = warp >> 2, wq = warp & 3;
            float run_m[NCH], run_l[NCH], acc[NCH][4][DPL];
#pragma unroll
            for (int ch = 0; ch < NCH; ch++) {
                run_m[ch] = -INFINITY;
                run_l[ch] = 0.0f;
#pragma unroll
                for (int h = 0; h < 4; h++)
#pragma unroll
                    for (int dd = 0; dd < DPL; dd++) acc[ch][h][dd] = 0.0f;
            }
            const int hsel = lane & 3, jsel = lane >> 2;
#pragma unroll 1
            for (int i = 0; i < np; i++) {
                const uint32_t aic = ai + i;
                if ((int)(aic % NGA) != ga) continue;
                const int s = aic % MG_KV_STAGES;
                mg_mbar_wait(&c.kv_full[s], (aic / MG_KV_STAGES) & 1, sp, 6);
                const __half* Kp = reinterpret_cast<const __half*>(stages + (s * 2 + 0) * PAGE_BYTES);
                const __half* Vp = reinterpret_cast<const __half*>(stages + (s * 2 + 1) * PAGE_BYTES);
                const int ntok = min(MG_PAGE, T - (pg0 + i) * MG_PAGE);
#pragma unroll
                for (int ch = 0; ch < NCH; ch++) {
                    float qv[4][DPL];
#pragma unroll
                    for (int h = 0; h < 4; h++)
#pragma unroll
                        for (int dd = 0; dd < DPL; dd++) qv[h][dd] = q_s[(ch * 4 + h) * HS + lane * DPL + dd];
                    float sc[32];
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const int tk = wq * 8 + j;
                        float kf[DPL];
#pragma unroll
                        for (int dd = 0; dd < DPL; dd++) kf[dd] = __half2float(Kp[tk * HS + lane * DPL + dd]);
#pragma unroll
                        for (int h = 0; h < 4; h++) {
                            float x = 0.0f;
#pragma unroll
                            for (int dd = 0; dd < DPL; dd++) x += qv[h][dd] * kf[dd];
                            sc[j * 4 + h] = x;
                        }
                    }
                    float sv = warp_transpose_reduce32_m(sc, lane);      // score of (token jsel, head hsel)
                    const bool valid = (wq * 8 + jsel < ntok) && (ch * 4 + hsel < G);
                    sv = valid ? sv : -INFINITY;
                    float mx = sv;
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
                    const float m_new = fmaxf(run_m[ch], mx);
                    const float m_use = (m_new == -INFINITY) ? 0.0f : m_new;
                    const float corr = __expf(run_m[ch] - m_use);
                    const float p = valid ? __expf(sv - m_use) : 0.0f;
                    float ps = p;
                    ps += __shfl_xor_sync(0xffffffffu, ps, 4);
                    ps += __shfl_xor_sync(0xffffffffu, ps, 8);
                    ps += __shfl_xor_sync(0xffffffffu, ps, 16);
                    run_l[ch] = run_l[ch] * corr + ps;
                    run_m[ch] = m_new;
#pragma unroll
                    for (int h = 0; h < 4; h++) {
                        const float chf = __shfl_sync(0xffffffffu, corr, h);
#pragma unroll
                        for (int dd = 0; dd < DPL; dd++) acc[ch][h][dd] *= chf;
                    }
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const int tk = wq * 8 + j;
                        float vf[DPL];
#pragma unroll
                        for (int dd = 0; dd < DPL; dd++) vf[dd] = (tk < ntok) ? __half2float(Vp[tk * HS + lane * DPL + dd]) : 0.0f;
#pragma unroll
                        for (int h = 0; h < 4; h++) {
                            const float ph = __shfl_sync(0xffffffffu, p, j * 4 + h);
#pragma unroll
                            for (int dd = 0; dd < DPL; dd++) acc[ch][h][dd] += ph * vf[dd];
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&c.kv_empty[s]);
            }
#pragma unroll
            for (int ch = 0; ch < NCH; ch++) {
#pragma unroll
                for (int h = 0; h < 4; h++)
#pragma unroll
                    for (int dd = 0; dd < DPL; dd++) w_acc[(warp * GMAX + ch * 4 + h) * HS + lane * DPL + dd] = acc[ch][h][dd];
                if (lane < 4) {
                    w_ml[(warp * GMAX + ch * 4 + lane) * 2 + 0] = run_m[ch];
                    w_ml[(warp * GMAX + ch * 4 + lane) * 2 + 1] = run_l[ch];
                }
            }
        }
        ai += np;
        cons_bar();
        // CTA-level merge of the AW warps -> split partial (unnormalised): thread = (head, dim)
        for (int i = tid; i < G * HS; i += MG_CONS) {
            const int h = i / HS, dd = i - h * HS;
            float Mx = -INFINITY;
#pragma unroll
            for (int w = 0; w < AW; w++) Mx = fmaxf(Mx, w_ml[(w * GMAX + h) * 2]);
            const float Mu = (Mx == -INFINITY) ? 0.0f : Mx;
            float Ls = 0.0f, O = 0.0f;
#pragma unroll
            for (int w = 0; w < AW; w++) {
                const float e = __expf(w_ml[(w * GMAX + h) * 2] - Mu);
                Ls += w_ml[(w * GMAX + h) * 2 + 1] * e;
                O += w_acc[(w * GMAX + h) * HS + dd] * e;
            }
            const size_t pidx = ((size_t)m * a.n_heads + (kvh * G + h)) * a.max_splits + sp_i;
            a.att_o[pidx * HS + dd] = O;
            if (dd == 0) {
                a.att_ml[pidx * 2 + 0] = Mx;
                a.att_ml[pidx * 2 + 1] = Ls;
            }
        }
        cons_bar();
    }
    ai_io = ai;
}

// long contexts (> 4 splits): one warp per (row, head) folds the splits into split slot 0, normalised
__device__ __noinline__ void attention_merge_coop(const DlDecodeArgs& a) {
    const MgCtx c;
    const int hs = a.head_size, warp = c.tid >> 5, lane = c.tid & 31;
    const int n_items = a.M * a.n_heads;
#pragma unroll 1
    for (int item = c.cta * (MG_CONS / 32) + warp; item < n_items; item += c.G * (MG_CONS / 32)) {
        const int m = item / a.n_heads;
        const int n_eff = mg_n_eff(a.pos[m], a.att_splits);
        const size_t base = (size_t)item * a.max_splits;
        float ms = -INFINITY, ls = 0.0f;
        if (lane < n_eff) {
            ms = __ldcg(&a.att_ml[(base + lane) * 2]);
            ls = __ldcg(&a.att_ml[(base + lane) * 2 + 1]);
        }
        const float Mx = warp_max(ms);
        const float Mu = (Mx == -INFINITY) ? 0.0f : Mx;
        const float e = (lane < n_eff) ? __expf(ms - Mu) : 0.0f;
        const float Ls = warp_sum(ls * e);
        const float inv = (Ls > 0.0f) ? 1.0f / Ls : 0.0f;
        for (int d0 = lane; d0 < hs; d0 += 32) {
            float O = 0.0f;
            for (int s2 = 0; s2 < n_eff; s2++) O += __ldcg(&a.att_o[(base + s2) * hs + d0]) * __shfl_sync(0xffffffffu, e, s2);
            a.att_o[base * hs + d0] = O * inv;     // slot 0 of this (row, head) is only read by this lane, above
        }
        __syncwarp();
        if (lane == 0) {
            a.att_ml[base * 2 + 0] = 0.0f;
            a.att_ml[base * 2 + 1] = 1.0f;
        }
    }
}

// =================================================================================================
// the kernel
// =================================================================================================
template <int MT>
__global__ void __launch_bounds__(MG_THREADS, 1) decode_mega_kernel(const __grid_constant__ DlDecodeArgs a) {
    const MgCtx c;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < MG_NS; s++) {
            mbar_init(&c.full[s], 1);
            mbar_init(&c.empty[s], 4);
        }
        for (int s = 0; s < MG_KV_STAGES; s++) {
            mbar_init(&c.kv_full[s], 1);
            mbar_init(&c.kv_empty[s], 4);
        }
        mbar_fence_init();
    }
    if (tid < MG_MAX_M) {                                    // running greedy best of this CTA
        float* run = c.misc + MG_BEST_OFF + 4 * MG_MAX_M * 2;
        run[tid * 2] = -INFINITY;
        run[tid * 2 + 1] = __int_as_float(0x7fffffff);
    }
    __syncthreads();
    const uint32_t step = ld_volatile_u32(a.step_counter) + 1u;          // nobody writes it before the last grid barrier

    if (warp == 16) { weight_producer(a); return; }
    if (warp == 17) { kv_producer(a); return; }

    // ------------------------------------------ consumers ------------------------------------------
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int M = a.M, L = a.n_layers, N = a.world;
    const bool multi = N > 1;
    const uint32_t ebase = step * (uint32_t)(2 * L);       // LL epochs of this step: ebase + 2l + 1 (wo), ebase + 2l + 2 (w2)
    uint32_t it = 0, ai = 0;
    int cur = 0;
    unsigned long long* tr = (a.trace != nullptr && c.cta == 0 && tid == 0) ? a.trace : nullptr;
    // do any rows need the cooperative split merge?  (uniform over the grid: same inputs everywhere)
    bool coop = false;
    for (int m = 0; m < M; m++) coop = coop || (mg_n_eff(a.pos[m], a.att_splits) > MG_SMALL_SPLITS);
    const int att_sel = a.head_size * 8 + (a.n_heads / a.n_kv_heads > 4 ? 2 : 1);
    int R0, R1;

#pragma unroll 1
    for (int l = 0; l < L; l++) {
        if (tr) tr[l * 8 + 0] = globaltimer_ns();
        // ---- attention block: norm -> qkv ----
        {
            mg_cta_rows(a.q_dim + 2 * a.kv_dim, c.G, 2, c.cta, R0, R1);
            const bool merge = multi && l > 0;
            // with a replicated residual (world > 1) CTA 0 materialises x of this layer
            float* xo = (multi && c.cta == 0) ? a.x[l == 0 ? 0 : (cur ^ 1)] : nullptr;
            if (R1 > R0 || xo != nullptr)
                prologue_norm(a, l == 0 ? nullptr : a.x[cur], xo, merge ? a.ll_ffn[a.rank] : nullptr,
                              ebase + 2 * l /* w2 of layer l-1 */, a.layers[l].rms_att);
            if (merge) cur ^= 1;
        }
        if (tr) tr[l * 8 + 1] = globaltimer_ns();
        gemv_phase<MT>(a, PH_QKV, l, it, 0u);
        grid_sync(a.grid_bar, c.G, sp);
        if (tr) tr[l * 8 + 2] = globaltimer_ns();
        // ---- attention ----
        cons_kv_bar();
        switch (att_sel) {
            case 32 * 8 + 1: attention_phase<32, 1>(a, ai); break;
            case 32 * 8 + 2: attention_phase<32, 2>(a, ai); break;
            case 64 * 8 + 1: attention_phase<64, 1>(a, ai); break;
            case 64 * 8 + 2: attention_phase<64, 2>(a, ai); break;
            case 128 * 8 + 1: attention_phase<128, 1>(a, ai); break;
            default: attention_phase<128, 2>(a, ai); break;
        }
        grid_sync(a.grid_bar, c.G, sp);
        if (coop) {
            attention_merge_coop(a);
            grid_sync(a.grid_bar, c.G, sp);
        }
        if (tr) tr[l * 8 + 3] = globaltimer_ns();
        // ---- wo ----
        mg_cta_rows(a.dim, c.G, 32, c.cta, R0, R1);
        if (R1 > R0) prologue_att(a, coop);
        gemv_phase<MT>(a, PH_WO, l, it, ebase + 2 * l + 1);
        if (!multi) grid_sync(a.grid_bar, c.G, sp);
        if (tr) tr[l * 8 + 4] = globaltimer_ns();
        // ---- feed-forward block: norm -> w1|w3 ----
        {
            mg_cta_rows(2 * a.hid, c.G, 2, c.cta, R0, R1);
            float* xo = (multi && c.cta == 0) ? a.x[cur ^ 1] : nullptr;
            if (R1 > R0 || xo != nullptr)
                prologue_norm(a, a.x[cur], xo, multi ? a.ll_att[a.rank] : nullptr, ebase + 2 * l + 1, a.layers[l].rms_ffn);
            if (multi) cur ^= 1;
        }
        if (tr) tr[l * 8 + 5] = globaltimer_ns();
        gemv_phase<MT>(a, PH_W13, l, it, 0u);
        grid_sync(a.grid_bar, c.G, sp);
        if (tr) tr[l * 8 + 6] = globaltimer_ns();
        // ---- w2 ----
        mg_cta_rows(a.dim, c.G, 32, c.cta, R0, R1);
        if (R1 > R0) prologue_hidden(a);
        gemv_phase<MT>(a, PH_W2, l, it, ebase + 2 * l + 2);
        if (!multi) grid_sync(a.grid_bar, c.G, sp);
        if (tr) tr[l * 8 + 7] = globaltimer_ns();
    }

    // ---- final norm -> logits (OP_MATMUL wcls + logits gather, src/llm.cpp:394-432) ----
    mg_cta_rows(a.vocab_loc, c.G, 2, c.cta, R0, R1);
    if (R1 > R0) prologue_norm(a, a.x[cur], nullptr, multi ? a.ll_ffn[a.rank] : nullptr, ebase + 2 * L, a.rms_final);
    gemv_phase<MT>(a, PH_CLS, 0, it, 0u);
    if (tr) tr[L * 8 + 0] = globaltimer_ns();
    if (a.greedy) {
        // CTA candidates -> global; CTA 0 reduces them, exchanges the per-rank winners as LL words, picks the token
        if (tid < M) {
            const float* run = c.misc + MG_BEST_OFF + 4 * MG_MAX_M * 2;
            a.cta_best[((size_t)c.cta * MG_MAX_M + tid) * 2] = run[tid * 2];
            a.cta_best[((size_t)c.cta * MG_MAX_M + tid) * 2 + 1] = run[tid * 2 + 1];
        }
    } else if (multi) {
        cons_bar();
        if (tid == 0) __threadfence_system();      // this CTA's logits stores to the peers, before the grid barrier
    }
    grid_sync(a.grid_bar, c.G, sp);
    if (c.cta != 0) return;
    if (a.greedy) {
        if (warp < M && warp >= a.row0) {
            const int m = warp, lane = tid & 31;
            float bv = -INFINITY;
            int bi = 0x7fffffff;
            for (int i = lane; i < c.G; i += 32)
                argmax_combine2(bv, bi, __ldcg(&a.cta_best[((size_t)i * MG_MAX_M + m) * 2]),
                                __float_as_int(__ldcg(&a.cta_best[((size_t)i * MG_MAX_M + m) * 2 + 1])));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                argmax_combine2(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
            if (multi) {
                // rank winners travel as two LL words per (source rank, row); every rank picks the same token
                if (lane < 2 * N) {
                    const int p = lane >> 1, wsel = lane & 1;
                    const uint32_t payload = wsel ? (uint32_t)bi : __float_as_uint(bv);
                    st_volatile_u64(reinterpret_cast<unsigned long long*>(a.ll_tok[p]) + ((size_t)a.rank * MG_MAX_M + m) * 2 + wsel,
                                    ((unsigned long long)step << 32) | payload);
                }
                float ov = -INFINITY;
                int oi = 0x7fffffff;
                if (lane < N) {
                    const unsigned long long* wp = reinterpret_cast<const unsigned long long*>(a.ll_tok[a.rank]) + ((size_t)lane * MG_MAX_M + m) * 2;
                    unsigned long long w0, w1;
                    do {
                        w0 = ld_volatile_u64(wp);
                        w1 = ld_volatile_u64(wp + 1);
                        sp.tick(7);
                    } while ((uint32_t)(w0 >> 32) != step || (uint32_t)(w1 >> 32) != step);
                    ov = __uint_as_float((uint32_t)w0);
                    oi = (int)(uint32_t)w1;
                }
                bv = ov; bi = oi;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1)
                    argmax_combine2(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
            }
            if (lane == 0) {
                if (bi == 0x7fffffff) bi = 0;
                a.tokens_out[m - a.row0] = bi;
                if (a.advance) { a.tokens[m] = bi; a.pos[m] += 1; }
                if (a.history != nullptr) a.history[(size_t)((step - 1u) % (uint32_t)a.history_cap) * M + m] = bi;
            }
        }
    } else if (multi) {
        if (tid < N) st_release_sys(a.flag_logits[tid], step);
    }
    cons_bar();
    if (tid == 0) {
        *a.step_counter = step;
        if (tr) tr[L * 8 + 1] = globaltimer_ns();
    }
}

// =================================================================================================
// host side
// =================================================================================================
extern "C" int dl_decode_mega_plan(int d, int n, int align, int grid, int cta, int* r0, int* r1, int* n_chunks) {
    int R0, R1;
    mg_cta_rows(d, grid, align, cta, R0, R1);
    int chunks = 0;
    const int KT = n / 128;
    for (int r = R0; r < R1;) {
        const int nr = (128 - (r & 127)) < (R1 - r) ? (128 - (r & 127)) : (R1 - r);
        const int kp = mg_kpack(nr);
        chunks += (KT + kp - 1) / kp;
        r += nr;
    }
    *r0 = R0; *r1 = R1; *n_chunks = chunks;
    return 0;
}

extern "C" int dl_decode_mega_supported(const DlDecodeArgs* a, int grid) {
    if (a->M < 1 || a->M > MG_MAX_M) { dl_set_error("decode_mega: M=%d outside 1..%d", a->M, MG_MAX_M); return -1; }
    if (a->row0 < 0 || a->row0 >= a->M) { dl_set_error("decode_mega: row0=%d", a->row0); return -1; }
    if (a->head_size != 32 && a->head_size != 64 && a->head_size != 128) { dl_set_error("decode_mega: head_size %d", a->head_size); return -1; }
    if (a->n_kv_heads < 1 || a->n_heads % a->n_kv_heads != 0 || a->n_heads / a->n_kv_heads > 8) { dl_set_error("decode_mega: %d query heads on %d kv heads (max 8 per kv head)", a->n_heads, a->n_kv_heads); return -1; }
    if (a->dim % 128 || a->q_dim % 128 || a->hid % 128) { dl_set_error("decode_mega: dim/q_dim/hid must be multiples of 128"); return -1; }
    if ((a->q_dim + 2 * a->kv_dim) % 2 || a->vocab_loc % 2 || a->kv_dim % 2 || a->head_size % 2) { dl_set_error("decode_mega: odd row counts"); return -1; }
    // prologue_norm parks the merged row (f32) behind its Q80 slabs
    if ((long long)a->M * a->dim * 4 + (long long)(a->dim / 128) * a->M * XQ_ROW_BYTES > MG_ACT_BYTES) { dl_set_error("decode_mega: M*dim=%lld does not fit the shared-memory row buffer", (long long)a->M * a->dim); return -1; }
    const long long kmax = a->hid > a->dim ? (a->hid > a->q_dim ? a->hid : a->q_dim) : (a->dim > a->q_dim ? a->dim : a->q_dim);
    if (kmax / 128 * a->M * XQ_ROW_BYTES > MG_ACT_BYTES) { dl_set_error("decode_mega: activation slabs (%lld B) exceed %d B of shared memory", kmax / 128 * a->M * XQ_ROW_BYTES, MG_ACT_BYTES); return -1; }
    if (a->world < 1 || a->world > DL_MAX_PEERS || a->rank < 0 || a->rank >= a->world) { dl_set_error("decode_mega: rank/world"); return -1; }
    if (a->world > 1 && !a->q80_sync) { dl_set_error("decode_mega: world > 1 ships Q80 partial sums (q80_sync must be on)"); return -1; }
    if (a->att_splits < 1 || a->att_splits > a->max_splits || a->max_splits > 32) { dl_set_error("decode_mega: att_splits=%d max_splits=%d", a->att_splits, a->max_splits); return -1; }
    if (grid < 1) { dl_set_error("decode_mega: grid=%d", grid); return -1; }
    return 0;
}

template <int MT>
static int launch_mega(const DlDecodeArgs& a, int grid, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(decode_mega_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL);
        if (e != cudaSuccess) { dl_set_error("decode_mega smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(MG_THREADS);
    cfg.dynamicSmemBytes = SM_TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident: the grid barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, decode_mega_kernel<MT>, a);
    if (e != cudaSuccess) { dl_set_error("decode_mega launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

extern "C" int dl_decode_mega(const DlDecodeArgs* args, int grid, void* stream) {
    if (dl_decode_mega_supported(args, grid) != 0) return -1;
    static_assert(SM_TOTAL <= 227 * 1024, "shared memory map exceeds 227 KB");
    cudaStream_t st = (cudaStream_t)stream;
    if (args->M == 1) return launch_mega<1>(*args, grid, st);
    if (args->M == 2) return launch_mega<2>(*args, grid, st);
    return launch_mega<4>(*args, grid, st);
}
