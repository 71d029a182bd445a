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
#define MG_NPROD 3                        // weight producer warps (16..18): chunk `it` belongs to producer it % 3
#define MG_KV_WARP (16 + MG_NPROD)        // KV page producer warp
#define MG_THREADS (MG_CONS + 32 * MG_NPROD + 32)   // 20 warps = 5 per scheduler: still 96 registers per thread
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
#define SM_MISC (SM_QS + 8 * 128 * 4)                     // ssq / argmax / w_ml scratch (2.5 KB)
#define SM_ROWS (SM_MISC + 2560)                          // per-row constants of this launch (MgRowInfo[4]) + RoPE table
#define SM_ROPE (SM_ROWS + 256)                           // [4 rows][64 pairs] (cos, sin)
#define SM_ARGS (SM_ROPE + MG_MAX_M * 64 * 8)             // copy of the kernel arguments (see mg_args())
#define SM_HBAR (SM_ARGS + 1024)                          // mbarrier of the hidden-activation bulk copy
#define SM_TOTAL (SM_HBAR + 16)
static_assert(sizeof(DlDecodeLayer) == 64, "layer table entry");
static_assert(sizeof(DlDecodeArgs) <= 1024, "DlDecodeArgs outgrew its shared-memory copy");

extern void dl_set_error(const char* fmt, ...);

// Spin guard: every wait loop ticks a per-thread counter; every 2^16 ticks it compares the global timer with the
// deadline of the launch (kernel start + timeout_ns, kept in shared memory) and traps when it has passed -- the launch dies
// with an error instead of hanging the GPU.  `code` says which wait it was (left in mapped host memory, see DlDecodeArgs).
__device__ __noinline__ void mg_deadline_check(int code);
struct MgSpin {
    unsigned n;
    DL_DEVINL void init(unsigned long long, uint32_t*) { n = 0; }
    DL_DEVINL void tick(int code) {
        if (((++n) & 0xffffu) == 0u) mg_deadline_check(code);
    }
    DL_DEVINL void reset() {}
};

DL_DEVINL void mg_mbar_wait(uint64_t* bar, uint32_t parity, MgSpin& sp, int code) {
    while (!mbar_try_wait(bar, parity)) sp.tick(code);
    sp.reset();
}
// for the producer warps: they share their schedulers with four consumer warps each, and a hot try_wait loop (210 M
// instructions per token, measured) takes issue slots from them
DL_DEVINL void mg_mbar_wait_idle(uint64_t* bar, uint32_t parity, MgSpin& sp, int code) {
    while (!mbar_try_wait(bar, parity)) {
        __nanosleep(200);
        sp.tick(code);
    }
}
DL_DEVINL void cons_bar() { asm volatile("bar.sync 1, %0;" ::"n"(MG_CONS) : "memory"); }
DL_DEVINL void cons_kv_bar() { asm volatile("bar.sync 2, %0;" ::"n"(MG_CONS + 32) : "memory"); }
DL_DEVINL void cons_kv_bar2() { asm volatile("bar.sync 3, %0;" ::"n"(MG_CONS + 32) : "memory"); }
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

// Grid barrier among the consumer threads of all CTAs (all CTAs are co-resident: cooperative launch).  The counter
// starts at 0 in every launch (the last CTA to leave the kernel clears it), so the k-th barrier completes at k * G:
// one fire-and-forget release-add per CTA, then an acquire spin -- no ticket round trip.
DL_DEVINL void red_release_gpu_add(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DL_DEVINL unsigned ld_acquire_gpu_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
DL_DEVINL void grid_sync(unsigned* bar, unsigned G, unsigned& k, MgSpin& sp) {
    k++;
    cons_bar();
    if (threadIdx.x == 0) {
        red_release_gpu_add(bar, 1u);
        const unsigned target = k * G;
        while (ld_acquire_gpu_u32(bar) < target) sp.tick(1);
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
// what the phases need to know about a batch row, fetched ONCE per launch: positions, slots and block tables do not
// change inside a step, and every global load after a grid barrier is an L2 round trip shared with 147 other CTAs
struct MgRowInfo {
    int pos, page, n_eff, token, slot, pad0, pad1, pad2;   // position, KV page of that position, attention splits in use, input token
};
struct MgCtx {
    uint64_t *full, *empty, *kv_full, *kv_empty;
    uint8_t *ring, *act;
    float *red, *out_s, *q_s, *misc;
    struct MgRowInfo* rows;
    DlDecodeLayer* lay;            // [2]: the layer table entries of layers l and l + 1 (slot l & 1), see the kernel
    float2* rope;
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
        rows = reinterpret_cast<struct MgRowInfo*>(mg_smem + SM_ROWS);
        lay = reinterpret_cast<DlDecodeLayer*>(mg_smem + SM_ROWS + 128);
        rope = reinterpret_cast<float2*>(mg_smem + SM_ROPE);
        tid = threadIdx.x;
        cta = blockIdx.x;
        G = gridDim.x;
    }
};

// The phase functions are real calls (one copy of each in the instruction stream).  Reaching the kernel parameters
// through a reference from there means generic loads from parameter space, which miss after every acquire of a grid
// barrier (~0.8 us per phase entry, measured); a copy in shared memory costs one LDS instead.
DL_DEVINL const DlDecodeArgs& mg_args() { return *reinterpret_cast<const DlDecodeArgs*>(mg_smem + SM_ARGS); }

__device__ __noinline__ void mg_deadline_check(int code) {
    const DlDecodeArgs& a = mg_args();
    const unsigned long long deadline = *reinterpret_cast<const unsigned long long*>(mg_smem + SM_HBAR + 8);
    if (a.timeout_ns == 0ull || globaltimer_ns() < deadline) return;
    if (a.debug != nullptr) {
        asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(a.debug), "r"(0x80000000u | (uint32_t)code * 65536u | blockIdx.x) : "memory");
        __threadfence_system();
    }
    __trap();
}

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
// Walks this CTA's pieces (row slices of one Q40T row tile over the full K extent) through every phase of every layer,
// in the order the consumers eat them.
struct MgPieceCursor {
    int l, ph, r, R1;
    const uint8_t* tile0;          // first k-tile of the piece's row tile
    int r0, nr, kpack, nch, KT;
    uint32_t it0, it_next;         // global chunk index of the piece's first chunk / of the piece after it
    bool done;
    DL_DEVINL void init() { l = 0; ph = PH_QKV - 1; r = 0; R1 = 0; it0 = 0; it_next = 0; done = false; }
    DL_DEVINL bool next(const DlDecodeArgs& a, int cta, int G) {
        const uint8_t* w = nullptr;
        while (r >= R1) {                                   // next phase with rows for this CTA
            if (ph == PH_CLS) { done = true; return false; }
            ph++;
            if (ph > PH_W2 && l + 1 < a.n_layers) { l++; ph = PH_QKV; }
            const MgPhase p = mg_phase(a, a.layers + l, ph);
            int R0;
            mg_cta_rows(p.d, G, p.align, cta, R0, R1);
            r = R0;
            KT = p.n >> 7;
        }
        w = mg_phase(a, a.layers + l, ph).w;
        const int rt = r >> 7;
        r0 = r & 127;
        nr = min(128 - r0, R1 - r);
        kpack = mg_kpack(nr);
        nch = (KT + kpack - 1) / kpack;
        tile0 = w + (size_t)rt * KT * Q40T_TILE_BYTES;
        it0 = it_next;
        it_next = it0 + nch;
        r += nr;
        return true;
    }
};

DL_DEVINL void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}

// The three producer warps: warp `me` owns the stages (chunks) with it % 3 == me.  While its next ring slot is still
// occupied -- the consumers are in a prologue, the attention phase or a barrier and the ring (~2.5 us of the stream) is
// full -- the warp spends the time pulling ITS stages up to pf_dist ahead into L2 (cp.async.bulk.prefetch.L2), so HBM
// keeps streaming through the bubbles of the dependency chain and the next matmul phases are fed from L2.  Prefetches are
// only issued while blocked: they never compete with the ring's own loads for issue slots.
__device__ __noinline__ void weight_producer() {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const uint64_t pol = l2_policy_evict_first();
    const int lane = c.tid & 31;
    const uint32_t me = (uint32_t)((c.tid >> 5) - 16);
    MgPieceCursor ld, pf;
    ld.init();
    pf.init();
    // prefetch cursor: piece pf, my next stage pf_ci inside it (global index pf.it0 + pf_ci)
    int pf_ci = 0;
    bool pf_on = a.pf_dist > 0;
    auto pf_seek = [&]() {            // position pf_ci on my first stage of the current prefetch piece
        pf_ci = (int)((me + MG_NPROD - pf.it0 % MG_NPROD) % MG_NPROD);
    };
    auto pf_advance = [&]() {         // next stage of my share, crossing pieces; clears pf_on at the end of the schedule
        pf_ci += MG_NPROD;
        while (pf_on && pf_ci >= pf.nch) {
            if (!pf.next(a, c.cta, c.G)) { pf_on = false; break; }
            pf_seek();
        }
    };
    if (pf_on) {
        if (pf.next(a, c.cta, c.G)) { pf_seek(); pf_ci -= MG_NPROD; pf_advance(); } else pf_on = false;
    }
#pragma unroll 1
    while (ld.next(a, c.cta, c.G)) {
        int ci = (int)((me + MG_NPROD - ld.it0 % MG_NPROD) % MG_NPROD);
#pragma unroll 1
        for (; ci < ld.nch; ci += MG_NPROD) {
            const uint32_t itc = ld.it0 + ci;
            const int kt0 = ci * ld.kpack;
            const int nkt = min(ld.kpack, ld.KT - kt0);
            const int nr = ld.nr;
            const int s = itc % MG_NS;
            if (itc >= MG_NS) {
                const uint32_t par = ((itc / MG_NS) & 1) ^ 1;
                // Blocked on the consumers?  Use the time for the stages ahead.  The burst polls with test_wait:
                // try_wait suspends the warp for a hardware time slice, which leaves room for only a few prefetches
                // in a bubble of a few microseconds.
                while (pf_on && !mbar_test_wait(&c.empty[s], par)) {
                    while (pf_on && pf.it0 + (uint32_t)pf_ci <= itc) pf_advance();             // never behind the ring
                    if (!pf_on || pf.it0 + (uint32_t)pf_ci >= itc + (uint32_t)a.pf_dist) break;
                    const int pk0 = pf_ci * pf.kpack;
                    const int pn = min(pf.kpack, pf.KT - pk0);
                    if (lane < pn) {
                        const uint8_t* t = pf.tile0 + (size_t)(pk0 + lane) * Q40T_TILE_BYTES;
                        bulk_prefetch_l2(t + pf.r0 * 64, (uint32_t)(pf.nr * 64));
                        bulk_prefetch_l2(t + Q40T_SCALE_OFF + pf.r0 * 8, (uint32_t)(pf.nr * 8));
                    }
                    __syncwarp();
                    pf_advance();
                }
                mg_mbar_wait_idle(&c.empty[s], par, sp, 2);
            }
            uint8_t* st = c.ring + s * Q40T_TILE_BYTES;
            if (lane == 0) mbar_arrive_expect_tx(&c.full[s], (uint32_t)(nkt * nr * 72));
            __syncwarp();
            if (lane < nkt) {
                const uint8_t* t = ld.tile0 + (size_t)(kt0 + lane) * Q40T_TILE_BYTES;
                bulk_g2s_hint(st + lane * nr * 64, t + ld.r0 * 64, (uint32_t)(nr * 64), &c.full[s], pol);
            } else if (lane >= 16 && lane - 16 < nkt) {
                const int j = lane - 16;
                const uint8_t* t = ld.tile0 + (size_t)(kt0 + j) * Q40T_TILE_BYTES;
                bulk_g2s_hint(st + Q40T_NIB_BYTES + j * nr * 8, t + Q40T_SCALE_OFF + ld.r0 * 8, (uint32_t)(nr * 8), &c.full[s], pol);
            }
        }
    }
}

// =================================================================================================
// KV page producer warp: bulk copies of the K and V pages of this CTA's attention work items
// =================================================================================================
__device__ __noinline__ void kv_producer() {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int lane = c.tid & 31;
    const int page_bytes = MG_PAGE * a.head_size * 2;
    uint8_t* stages = c.act;
    const int n_items = a.M * a.n_kv_heads * a.att_splits;
    // Positions, slots and block tables are the same for every layer of a step: the page list of this CTA's first work
    // item (usually its only one) is read once, and only the page that receives this step's K/V row has to wait for the
    // grid barrier behind the qkv phase -- the older pages are in flight while the CTA still sits in that barrier.
    int f_m = 0, f_kvh = 0, f_pg0 = 0, f_pg1 = 0, f_tail = -1, f_page = 0;
    if (c.cta < n_items) {
        const int item = c.cta;
        const int sp_i = item % a.att_splits;
        f_kvh = (item / a.att_splits) % a.n_kv_heads;
        f_m = item / (a.att_splits * a.n_kv_heads);
        const int n_eff = c.rows[f_m].n_eff;
        if (sp_i < n_eff) {
            const int n_pages = c.rows[f_m].pos / MG_PAGE + 1;
            const int pps = (n_pages + n_eff - 1) / n_eff;
            f_pg0 = sp_i * pps;
            f_pg1 = min(n_pages, f_pg0 + pps);
            f_tail = n_pages - 1;
            if (f_pg0 + lane < f_pg1) f_page = a.block_table[(size_t)c.rows[f_m].slot * a.max_pages + f_pg0 + lane];
        }
    }
    uint32_t ai = 0;
    auto issue = [&](const uint8_t* kc, const uint8_t* vc, int page, int kvh) {     // lane 0 only
        const int s = ai % MG_KV_STAGES;
        if (ai >= MG_KV_STAGES) mg_mbar_wait_idle(&c.kv_empty[s], ((ai / MG_KV_STAGES) & 1) ^ 1, sp, 3);
        const size_t poff = ((size_t)page * a.n_kv_heads + kvh) * (size_t)page_bytes;
        mbar_arrive_expect_tx(&c.kv_full[s], 2 * page_bytes);
        bulk_g2s(stages + (s * 2 + 0) * page_bytes, kc + poff, page_bytes, &c.kv_full[s]);
        bulk_g2s(stages + (s * 2 + 1) * page_bytes, vc + poff, page_bytes, &c.kv_full[s]);
    };
#pragma unroll 1
    for (int l = 0; l < a.n_layers; l++) {
        const uint8_t* kc = reinterpret_cast<const uint8_t*>(a.layers[l].k_cache);
        const uint8_t* vc = reinterpret_cast<const uint8_t*>(a.layers[l].v_cache);
        cons_kv_bar();                 // (1) this CTA's qkv phase is over: the staging area is free
        // first item, pages that do not change in this step: at most one ring-full, so nothing here waits on a consumer
        int early = 0;
        {
            const int cnt = min(min(32, f_pg1 - f_pg0), MG_KV_STAGES);
            for (int j = 0; j < cnt; j++) {
                if (f_pg0 + j == f_tail) break;
                const int page = __shfl_sync(0xffffffffu, f_page, j);
                if (lane == 0) issue(kc, vc, page, f_kvh);
                ai++;
                early++;
                __syncwarp();
            }
        }
        cons_kv_bar2();                // (2) the grid barrier behind the KV writes has been passed
        fence_proxy_async_all();       // generic-proxy stores (this step's K/V row) -> async-proxy reads below
        for (int item = c.cta; item < n_items; item += c.G) {
            const int sp_i = item % a.att_splits;
            const int kvh = (item / a.att_splits) % a.n_kv_heads;
            const int m = item / (a.att_splits * a.n_kv_heads);
            const int pos = c.rows[m].pos;
            const int n_eff = c.rows[m].n_eff;
            if (sp_i >= n_eff) continue;
            const int n_pages = pos / MG_PAGE + 1;
            const int pps = (n_pages + n_eff - 1) / n_eff;
            const int pg0 = sp_i * pps, pg1 = min(n_pages, pg0 + pps);
            const int* bt = a.block_table + (size_t)c.rows[m].slot * a.max_pages;
            const bool first = item == c.cta;
            for (int base = pg0; base < pg1; base += 32) {
                const bool cached = first && base == pg0;                 // the window read before the layer loop
                const int my_page = cached ? f_page : ((base + lane < pg1) ? bt[base + lane] : 0);
                const int cnt = min(32, pg1 - base);
                for (int j = cached ? early : 0; j < cnt; j++) {
                    const int page = __shfl_sync(0xffffffffu, my_page, j);
                    if (lane == 0) issue(kc, vc, page, kvh);
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
DL_DEVINL float4 ll_dequant4(unsigned long long wd, unsigned long long ws) {
    const uint32_t p = (uint32_t)wd;
    const float d = __half2float(__ushort_as_half((unsigned short)(ws & 0xffffu)));
    return make_float4((float)(int8_t)(p & 0xff) * d, (float)(int8_t)((p >> 8) & 0xff) * d,
                       (float)(int8_t)((p >> 16) & 0xff) * d, (float)(int8_t)(p >> 24) * d);
}

// Q80-quantise one 32-value block held by ONE thread (8 float4) into the resident act slab.  One thread per block keeps
// the instruction count low: the prologues are replicated on every SM and were issue-bound with 8 lanes per block.
// (quantizeF32toQ80, src/nn/nn-quants.cpp:126-139: d = amax / 127, q = round-half-even(x / d), scale stored as fp16)
DL_DEVINL void q80_block_to_act(const float4 (&v)[8], uint8_t* act, int Ms, int m, int blk) {
    float amax = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; j++)
        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[j].x), fabsf(v[j].y)), fmaxf(fabsf(v[j].z), fabsf(v[j].w))));
    const float d = amax / 127.0f;
    const float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
    int sumq = 0;
    uint32_t pk[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int q0 = __float2int_rn(v[j].x * id), q1 = __float2int_rn(v[j].y * id);
        const int q2 = __float2int_rn(v[j].z * id), q3 = __float2int_rn(v[j].w * id);
        sumq += q0 + q1 + q2 + q3;
        pk[j] = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
    }
    uint8_t* rowp = act + ((size_t)(blk >> 2) * Ms + m) * XQ_ROW_BYTES;
    const int kb = blk & 3;
    uint4* qd = reinterpret_cast<uint4*>(rowp + kb * 32);
    qd[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    qd[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    reinterpret_cast<float*>(rowp + 128)[kb] = __half2float(__float2half_rn(d));
    reinterpret_cast<int*>(rowp + 144)[kb] = 8 * sumq;
}

// All CTAs read the same activation rows at the same moment; starting each CTA at a different offset spreads the requests
// over the L2 slices instead of queueing 148 of them on the same line.
// debug timeline inside the phase functions: slot 16 + k of the current layer (kept in shared memory by the kernel)
#define MG_ISTAMP(k)                                                                                          \
    do {                                                                                                      \
        if (a.trace != nullptr && blockIdx.x == a.trace_cta && threadIdx.x == 0)                              \
            a.trace[(size_t)c.rows[0].pad0 * 32 + 16 + (k)] = (unsigned long long)clock64();                               \
    } while (0)

DL_DEVINL int mg_rot(int cta, int total, int gran) {
    const int w = total / gran;
    return w > 0 ? ((cta * 5) % w) * gran : 0;
}
DL_DEVINL int mg_wrap(int i, int total) { return i >= total ? i - total : i; }

// ---- prologue 1: x = x_in + sum of the ranks' Q80 partial sums; RMSNorm; Q80 -> act slabs -------------------------
// (OP_MERGE_ADD + OP_INV_RMS + OP_RMS_NORM + OP_CAST, src/llm.cpp:204-239,317-342,394-419)
// Every CTA does the whole row (M*dim <= 16384 floats, L2 reads), so no CTA waits for another one.  Pass 1 (one float4
// per thread and step, loads batched: a dependent L2 round trip costs ~0.7 us here) merges, parks the row in shared
// memory behind the slabs and sums the squares; pass 2 (one 32-value block per thread) normalises and quantises.
// xsrc: residual rows ([M][dim] f32), or NULL = embedding rows of the batch tokens (layer 0: OP_EMBEDDING,
// src/llm.cpp:185-192).  Ms: row stride of the act slabs (the matmul phase's MT).
#define MG_PB 4       // float4 chunks per thread and batch in pass 1
#define MG_XS_STRIDE 9    // float4 per parked block (8 + 1 pad: lane-per-block reads stay bank-conflict free)
#define MG_MERGERS 16     // CTAs that reduce the ranks' partial sums (each 1/16 of the row) for everybody else
__device__ __noinline__ void prologue_norm(int Ms, const float* xsrc, float* x_out, const uint8_t* ll, uint32_t epoch,
                                           const float* norm_w) {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int M = a.M, dim = a.dim, tid = c.tid;
    const int cpr = dim >> 2;                              // float4 chunks per row
    const int total = M * cpr;
    const int nblk = dim >> 5;
    const int N = (ll != nullptr) ? a.world : 0;
    float4* xs = reinterpret_cast<float4*>(c.act + (size_t)(dim >> 7) * Ms * XQ_ROW_BYTES);     // [M * nblk][9] float4
    const int rot = mg_rot(c.cta, total, 32);              // slot sl of the sweep handles chunk (sl + rot) % total
    // model weights do not depend on the exchange: start them towards L1 now, pass 2 finds them there
    if (tid < nblk) asm volatile("prefetch.global.L1 [%0];" ::"l"(norm_w + tid * 32));
    MG_ISTAMP(0);
    if (N > 0) {
        // ---- merge-add, done ONCE per rank: polling the exchange words costs about a cycle per lane and word (8-byte
        // volatile loads), 16 k lane-loads per CTA with 8 ranks -- 10 us when every CTA did it for itself.  Merger CTA k
        // reduces blocks [k, k+1) * nb / K of the row(s), writes the merged residual to x_out and releases flag k.
        const int K = min(MG_MERGERS, c.G);
        unsigned* flags = a.grid_bar + 16;
        if (c.cta < K) {
            const int nb = M * nblk;
            const int b_lo = (int)((long long)nb * c.cta / K), b_hi = (int)((long long)nb * (c.cta + 1) / K);
            const size_t sstride = (size_t)MG_MAX_M * nblk * MG_LLW;
#pragma unroll 1
            for (int ch = b_lo * 8 + tid; ch < b_hi * 8; ch += MG_CONS) {
                const int m = ch / cpr, e4 = ch - m * cpr;
                float4 xv = (xsrc != nullptr) ? __ldcg(reinterpret_cast<const float4*>(xsrc + (size_t)m * dim) + e4)
                                              : reinterpret_cast<const float4*>(a.embedding + (size_t)c.rows[m].token * dim)[e4];
                // word (src, m, blk, w) lives at index ((src*4 + m)*nblk + blk)*9 + w; all sources are requested together
                const unsigned long long* base = reinterpret_cast<const unsigned long long*>(ll) + ((size_t)m * nblk + (e4 >> 3)) * MG_LLW;
                unsigned long long wd[DL_MAX_PEERS], ws[DL_MAX_PEERS];
                bool ok;
                do {
                    ok = true;
#pragma unroll
                    for (int sr = 0; sr < DL_MAX_PEERS; sr++) {
                        if (sr < N) {
                            wd[sr] = ld_volatile_u64(base + sr * sstride + (e4 & 7));
                            ws[sr] = ld_volatile_u64(base + sr * sstride + 8);
                        }
                    }
#pragma unroll
                    for (int sr = 0; sr < DL_MAX_PEERS; sr++)
                        if (sr < N) ok = ok && ((uint32_t)(wd[sr] >> 32) == epoch) && ((uint32_t)(ws[sr] >> 32) == epoch);
                    if (!ok) sp.tick(4);
                } while (!ok);
                // the same order on every rank: x + p0 + p1 + ... (bit-identical replicas of the residual stream)
#pragma unroll
                for (int sr = 0; sr < DL_MAX_PEERS; sr++) {
                    if (sr < N) {
                        const float4 p = ll_dequant4(wd[sr], ws[sr]);
                        xv.x += p.x; xv.y += p.y; xv.z += p.z; xv.w += p.w;
                    }
                }
                reinterpret_cast<float4*>(x_out + (size_t)m * dim)[e4] = xv;
            }
            cons_bar();
            if (tid == 0) {
                __threadfence();
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flags + c.cta), "r"(epoch) : "memory");
            }
        }
        if (tid < K) {
            while ((int)(ld_acquire_gpu_u32(flags + tid) - epoch) < 0) sp.tick(9);
        }
        cons_bar();
        xsrc = x_out;                                       // everybody reads the merged row
    }
    float ss[MG_MAX_M];
#pragma unroll
    for (int m = 0; m < MG_MAX_M; m++) ss[m] = 0.0f;
#pragma unroll 1
    for (int ch0 = tid; ch0 < total; ch0 += MG_PB * MG_CONS) {
        float4 x4[MG_PB];
#pragma unroll
        for (int u = 0; u < MG_PB; u++) {
            const int sl = ch0 + u * MG_CONS;
            x4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sl < total) {
                const int ch = mg_wrap(sl + rot, total);
                const int m = ch / cpr, e4 = ch - m * cpr;
                // the residual rows are rewritten by other CTAs every phase: read them from L2, never from a stale L1 line
                x4[u] = (xsrc != nullptr) ? __ldcg(reinterpret_cast<const float4*>(xsrc + (size_t)m * dim) + e4)
                                          : reinterpret_cast<const float4*>(a.embedding + (size_t)c.rows[m].token * dim)[e4];
            }
        }
#pragma unroll
        for (int u = 0; u < MG_PB; u++) {
            const int sl = ch0 + u * MG_CONS;
            if (sl < total) {
                const int ch = mg_wrap(sl + rot, total);
                const int m = ch / cpr, e4 = ch - m * cpr;
                if (N == 0 && x_out != nullptr) reinterpret_cast<float4*>(x_out + (size_t)m * dim)[e4] = x4[u];   // layer 0: CTA 0 keeps x
                xs[(size_t)(m * nblk + (e4 >> 3)) * MG_XS_STRIDE + (e4 & 7)] = x4[u];
                const float sq = x4[u].x * x4[u].x + x4[u].y * x4[u].y + x4[u].z * x4[u].z + x4[u].w * x4[u].w;
#pragma unroll
                for (int q = 0; q < MG_MAX_M; q++) ss[q] += (m == q) ? sq : 0.0f;
            }
        }
    }
    MG_ISTAMP(1);
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
    MG_ISTAMP(2);
#pragma unroll 1
    for (int b = tid; b < M * nblk; b += MG_CONS) {
        const int m = b / nblk, blk = b - m * nblk;
        float im = inv[0];
#pragma unroll
        for (int q = 1; q < MG_MAX_M; q++) im = (m == q) ? inv[q] : im;
        const float4* wn = reinterpret_cast<const float4*>(norm_w) + blk * 8;
        const float4* xb = xs + (size_t)b * MG_XS_STRIDE;
        float4 y[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const float4 w4 = wn[j], x4 = xb[j];
            y[j].x = w4.x * (im * x4.x); y[j].y = w4.y * (im * x4.y);
            y[j].z = w4.z * (im * x4.z); y[j].w = w4.w * (im * x4.w);
        }
        q80_block_to_act(y, c.act, Ms, m, blk);
    }
    cons_bar();
    MG_ISTAMP(3);
}

// ---- prologue 2: merge the attention splits, Q80 -> act slabs (K = q_dim)  (OP_CAST y->yq_slice, src/llm.cpp:296-301)
// pass 1: one float4 per thread (coalesced L2 reads of the split partials) -> merged, normalised output parked in shared
// memory; pass 2: one 32-value block per thread -> Q80.
__device__ __noinline__ void prologue_att(int Ms, bool merged) {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    const int M = a.M, cpr = a.q_dim >> 2, total = M * cpr, nblk = a.q_dim >> 5, hs = a.head_size, tid = c.tid;
    float4* xs = reinterpret_cast<float4*>(c.act + (size_t)(a.q_dim >> 7) * Ms * XQ_ROW_BYTES);     // [M * nblk][9] float4
    const int rot = mg_rot(c.cta, total, 8);
#pragma unroll 2
    for (int sl = tid; sl < total; sl += MG_CONS) {
        const int ch = mg_wrap(sl + rot, total);
        const int m = ch / cpr, e4 = ch - m * cpr;
        const int head = (e4 * 4) / hs, dd = (e4 * 4) - head * hs;
        const int n_eff = merged ? 1 : c.rows[m].n_eff;        // <= MG_SMALL_SPLITS here
        const size_t base = ((size_t)m * a.n_heads + head) * a.max_splits;
        float2 ml[MG_SMALL_SPLITS];
        float4 o[MG_SMALL_SPLITS];
#pragma unroll
        for (int sp = 0; sp < MG_SMALL_SPLITS; sp++) {
            ml[sp] = make_float2(-INFINITY, 0.0f);
            o[sp] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sp < n_eff) {
                ml[sp] = __ldcg(reinterpret_cast<const float2*>(a.att_ml + (base + sp) * 2));
                o[sp] = __ldcg(reinterpret_cast<const float4*>(a.att_o + (base + sp) * hs + dd));
            }
        }
        float Mx = -INFINITY;
#pragma unroll
        for (int sp = 0; sp < MG_SMALL_SPLITS; sp++) Mx = fmaxf(Mx, ml[sp].x);
        const float Mu = (Mx == -INFINITY) ? 0.0f : Mx;
        float L = 0.0f;
        float4 O = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int sp = 0; sp < MG_SMALL_SPLITS; sp++) {
            const float e = (sp < n_eff) ? __expf(ml[sp].x - Mu) : 0.0f;
            L += ml[sp].y * e;
            O.x += o[sp].x * e; O.y += o[sp].y * e; O.z += o[sp].z * e; O.w += o[sp].w * e;
        }
        const float il = (L > 0.0f) ? 1.0f / L : 0.0f;
        O.x *= il; O.y *= il; O.z *= il; O.w *= il;
        xs[(size_t)(m * nblk + (e4 >> 3)) * MG_XS_STRIDE + (e4 & 7)] = O;
    }
    cons_bar();
#pragma unroll 1
    for (int b = tid; b < M * nblk; b += MG_CONS) {
        const int m = b / nblk, blk = b - m * nblk;
        const float4* xb = xs + (size_t)b * MG_XS_STRIDE;
        float4 y[8];
#pragma unroll
        for (int j = 0; j < 8; j++) y[j] = xb[j];
        q80_block_to_act(y, c.act, Ms, m, blk);
    }
    cons_bar();
}

// ---- prologue 3: hidden activation f32 -> Q80 act slabs (K = hid)  (OP_CAST d->dq, src/llm.cpp:367-374)
// pass 1: coalesced float4 loads (a per-thread 128-byte stride over global memory costs 32 line requests per load:
// 3.6 us measured), parked in shared memory behind the slabs with one pad float4 per block; pass 2: one thread per
// 32-value block.  Plain loads on purpose: a bulk copy would need fence.proxy.async in every thread of the producing
// epilogue, which showed up as ~2 us of membar stalls per layer.
__device__ __noinline__ void prologue_hidden(int Ms) {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    const int M = a.M, nblk = a.hid >> 5, total = M * nblk, tid = c.tid;
    const size_t slab_bytes = (size_t)(a.hid >> 7) * Ms * XQ_ROW_BYTES;
    if (slab_bytes + (size_t)total * MG_XS_STRIDE * 16 <= MG_ACT_BYTES) {
        float4* xs = reinterpret_cast<float4*>(c.act + slab_bytes);               // [M * nblk][9] float4
        const int nch = total * 8;                                                // float4 chunks, rows are contiguous
        const int rot = mg_rot(c.cta, nch, 32);
        const float4* hp = reinterpret_cast<const float4*>(a.h);
#pragma unroll 1
        for (int ch0 = tid; ch0 < nch; ch0 += 8 * MG_CONS) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int sl = ch0 + u * MG_CONS;
                if (sl < nch) v[u] = __ldcg(hp + mg_wrap(sl + rot, nch));
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int sl = ch0 + u * MG_CONS;
                if (sl < nch) {
                    const int ch = mg_wrap(sl + rot, nch);
                    xs[(size_t)(ch >> 3) * MG_XS_STRIDE + (ch & 7)] = v[u];
                }
            }
        }
        cons_bar();
#pragma unroll 1
        for (int b = tid; b < total; b += MG_CONS) {
            const int m = b / nblk, blk = b - m * nblk;
            const float4* xb = xs + (size_t)b * MG_XS_STRIDE;
            float4 y[8];
#pragma unroll
            for (int j = 0; j < 8; j++) y[j] = xb[j];
            q80_block_to_act(y, c.act, Ms, m, blk);
        }
    } else {
        // rows too long for the staging area (several rows of a big model): per-thread global loads
#pragma unroll 1
        for (int b = tid; b < total; b += MG_CONS) {
            const int m = b / nblk, blk = b - m * nblk;
            const float4* hp = reinterpret_cast<const float4*>(a.h + (size_t)m * a.hid) + blk * 8;
            float4 h4[8];
#pragma unroll
            for (int j = 0; j < 8; j++) h4[j] = __ldcg(hp + j);
            q80_block_to_act(h4, c.act, Ms, m, blk);
        }
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
        const int p = c.rows[m].pos;
        if (region < 2) {
            const float2 csn = c.rope[m * 64 + (hd >> 1)];          // (cos, sin) of pos * freq, tabulated once per launch
            const float r0v = v0 * csn.x - v1 * csn.y, r1v = v0 * csn.y + v1 * csn.x;
            v0 = r0v; v1 = r1v;
        }
        if (region == 0) {
            *reinterpret_cast<float2*>(a.q + (size_t)m * a.q_dim + rrow) = make_float2(v0, v1);
        } else {
            const int page = c.rows[m].page;
            const size_t off = (((size_t)page * a.n_kv_heads + kvh) * MG_PAGE + (p % MG_PAGE)) * hs + hd;
            __half* base = reinterpret_cast<__half*>(region == 1 ? L->k_cache : L->v_cache);
            *reinterpret_cast<__half2*>(base + off) = __floats2half2_rn(v0, v1);
            fence_proxy_async_all();  // this KV row is read by bulk copies (async proxy) after the next grid barrier
        }
    }
}

// ---- epilogue: Q80 cast of a partial sum (32-row blocks) pushed as LL words into every rank's exchange buffer -------
// (OP_CAST y->ZQ slice + SYNC_NODE_SLICES, src/llm.cpp:308-314,381-387; the merge-add happens in the next prologue).
// One GPU uses the same hand-off (one source): the words themselves tell the next phase that the rows are there,
// which is cheaper than a grid barrier.
DL_DEVINL void epi_residual(const DlDecodeArgs& a, const MgCtx& c, int row, int nr, uint8_t* const* ll, uint32_t epoch) {
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

// unsigned bytes x signed bytes, 32-bit signed accumulate
DL_DEVINL int dp4a_us(uint32_t a_u8x4, int b_s8x4, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8x4), "r"(b_s8x4), "r"(c));
    return d;
}
DL_DEVINL void argmax_combine2(float& best, int& bi, float ov, int oi) {
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
}

// ---- epilogue: logits slice -> every rank, or (greedy) the running best of this CTA in shared memory ----------------
// (logits gather SYNC_NODE_SLICES_EXCEPT_ROOT, src/llm.cpp:426-432; Sampler::sample temperature == 0, src/tokenizer.cpp:472-502)
#define MG_BEST_OFF 256      // floats into misc: [4 warps][4][2] scratch, then [4][2] running best
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
__device__ __noinline__ uint32_t gemv_phase(int ph, int layer, uint32_t it_io, uint32_t epoch) {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const DlDecodeLayer* Lw = c.lay + (layer & 1);          // shared-memory copy (the weight pointers are not used here)
    const MgPhase P = mg_phase(a, Lw, ph);
    const int M = a.M;
    const int KT = P.n >> 7;
    const int tid = c.tid, g = tid >> 7, t = tid & 127, lane = tid & 31;
    uint32_t it = it_io;
    int R0, R1;
    mg_cta_rows(P.d, c.G, P.align, c.cta, R0, R1);
    int r = R0;
    if (ph == PH_QKV) MG_ISTAMP(4);
    if (ph == PH_W2) MG_ISTAMP(9);
#pragma unroll 1
    while (r < R1) {
        const int r0 = r & 127;
        const int nr = min(128 - r0, R1 - r);
        const int kpack = mg_kpack(nr);
        const int nch = (KT + kpack - 1) / kpack;
        // Thread -> work: the 128 threads of a group see a stage as up to 128 (k-tile, row) pairs; a thread owns one pair
        // and walks its 4 blocks.  The loop is bound by instruction issue (measured: its time tracks the instruction
        // count, not the memory pipeline), so everything here is about instructions per block: the activations of a
        // block are broadcast loads, the 4 scales of the row come with one 64-bit load, the high nibbles stay in place
        // (16 * nibble, unsigned bytes; the /16 is applied to their exact integer sum), the dot products accumulate
        // in the dp4a instructions themselves.
        const int j = t / nr, lr = t - j * nr;
        const bool mine = j < kpack;
        const int rot = (r0 + lr) >> 1;
        const int pofs = (mine ? t : 0);                           // pair slot inside a stage (idle threads alias pair 0)
        float acc[MT];
#pragma unroll
        for (int m = 0; m < MT; m++) acc[m] = 0.0f;
        // this group's stages of the piece: it + ci == g (mod 4)   (no loop over the other groups' stages: the skipped
        // iterations cost more in branch-resolving stalls than a third of the dot-product body, measured with ncu)
#pragma unroll 1
        for (int ci = (int)(((uint32_t)g - it) & 3u); ci < nch; ci += 4) {
            const uint32_t itc = it + ci;
            const int s = itc % MG_NS;
            if (!(a.dbg & 2)) mg_mbar_wait(&c.full[s], (itc / MG_NS) & 1, sp, 5);       // dbg 2: do not wait for the data (timing probe)
            const uint8_t* st = c.ring + s * Q40T_TILE_BYTES;
            const int kt = min(ci * kpack + j, KT - 1);            // a short last stage leaves stale slots: computed, dropped
            const bool live = mine && (ci * kpack + j < KT);
            if (!(a.dbg & 1)) {
                const uint8_t* nibp = st + pofs * 64;
                const uint2 sc2 = *reinterpret_cast<const uint2*>(st + Q40T_NIB_BYTES + pofs * 8);      // 4 fp16 scales
                const uint8_t* actp = c.act + (size_t)kt * MT * XQ_ROW_BYTES;     // slabs carry MT rows: no row test
                float accl[MT];
#pragma unroll
                for (int m = 0; m < MT; m++) accl[m] = 0.0f;
#pragma unroll
                for (int kb = 0; kb < 4; kb++) {
                    const uint4 nb = *reinterpret_cast<const uint4*>(nibp + (((kb + rot) & 3) << 4));
                    const uint32_t sw = (kb < 2) ? sc2.x : sc2.y;
                    const float dw = __half2float(__ushort_as_half((unsigned short)((kb & 1) ? (sw >> 16) : (sw & 0xffffu))));
                    const uint32_t lo0 = nb.x & 0x0F0F0F0Fu, hi0 = nb.x & 0xF0F0F0F0u;
                    const uint32_t lo1 = nb.y & 0x0F0F0F0Fu, hi1 = nb.y & 0xF0F0F0F0u;
                    const uint32_t lo2 = nb.z & 0x0F0F0F0Fu, hi2 = nb.z & 0xF0F0F0F0u;
                    const uint32_t lo3 = nb.w & 0x0F0F0F0Fu, hi3 = nb.w & 0xF0F0F0F0u;
#pragma unroll
                    for (int m = 0; m < MT; m++) {
                        const uint8_t* xr = actp + m * XQ_ROW_BYTES;
                        const int4 xl = *reinterpret_cast<const int4*>(xr + kb * 32);
                        const int4 xh = *reinterpret_cast<const int4*>(xr + kb * 32 + 16);
                        int is0 = dp4a_us(lo0, xl.x, 0), is1 = dp4a_us(hi0, xh.x, 0);
                        is0 = dp4a_us(lo1, xl.y, is0); is1 = dp4a_us(hi1, xh.y, is1);
                        is0 = dp4a_us(lo2, xl.z, is0); is1 = dp4a_us(hi2, xh.z, is1);
                        is0 = dp4a_us(lo3, xl.w, is0); is1 = dp4a_us(hi3, xh.w, is1);
                        const float dx = reinterpret_cast<const float*>(xr + 128)[kb];
                        const int s8 = reinterpret_cast<const int*>(xr + 144)[kb];
                        accl[m] += (float)(is0 + (is1 >> 4) - s8) * (dw * dx);
                    }
                }
#pragma unroll
                for (int m = 0; m < MT; m++) acc[m] = live ? acc[m] + accl[m] : acc[m];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&c.empty[s]);
        }
        it += nch;
        if (ph == PH_QKV) MG_ISTAMP(5);
        if (ph == PH_W2) MG_ISTAMP(10);
        if (mine) {
#pragma unroll
            for (int m = 0; m < MT; m++) c.red[((g * kpack + j) * MT + m) * nr + lr] = acc[m];
        }
        cons_bar();
        // out_s[m][row] = sum of the (group, k-tile) partials: four lanes per output (one per group), fixed order
#pragma unroll 1
        for (int base = 0; base < MT * nr * 4; base += MG_CONS) {
            const int o4 = base + tid;
            const bool live = o4 < MT * nr * 4;
            const int o = live ? (o4 >> 2) : 0, gg = o4 & 3;
            const int m = o / nr, l2 = o - m * nr;
            float vs = 0.0f;
            if (live) {
                for (int q = 0; q < kpack; q++) vs += c.red[((gg * kpack + q) * MT + m) * nr + l2];
            }
            vs += __shfl_xor_sync(0xffffffffu, vs, 1);
            vs += __shfl_xor_sync(0xffffffffu, vs, 2);
            if (live && gg == 0) c.out_s[m * 128 + l2] = vs;
        }
        cons_bar();
        if (ph == PH_QKV) MG_ISTAMP(6);
        if (ph == PH_W2) MG_ISTAMP(11);
        switch (ph) {
            case PH_QKV: epi_qkv(a, c, Lw, r, nr); break;
            case PH_WO: epi_residual(a, c, r, nr, a.ll_att, epoch); break;
            case PH_W13: epi_silu(a, c, r, nr); break;
            case PH_W2: epi_residual(a, c, r, nr, a.ll_ffn, epoch); break;
            default: epi_logits(a, c, r, nr); break;
        }
        if (ph == PH_QKV) MG_ISTAMP(7);
        if (ph == PH_W2) MG_ISTAMP(12);
        cons_bar();
        if (ph == PH_QKV) MG_ISTAMP(8);
        if (ph == PH_W2) MG_ISTAMP(13);
        r += nr;
    }
    return it;
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
__device__ __noinline__ uint32_t attention_phase(uint32_t ai_io) {
    const DlDecodeArgs& a = mg_args();
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
        const int pos = c.rows[m].pos;
        const int n_eff = c.rows[m].n_eff;
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
    return ai;
}

// long contexts (> 4 splits): one warp per (row, head) folds the splits into split slot 0, normalised
__device__ __noinline__ void attention_merge_coop() {
    const DlDecodeArgs& a = mg_args();
    const MgCtx c;
    const int hs = a.head_size, warp = c.tid >> 5, lane = c.tid & 31;
    const int n_items = a.M * a.n_heads;
#pragma unroll 1
    for (int item = c.cta * (MG_CONS / 32) + warp; item < n_items; item += c.G * (MG_CONS / 32)) {
        const int m = item / a.n_heads;
        const int n_eff = c.rows[m].n_eff;
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
        mbar_init(reinterpret_cast<uint64_t*>(mg_smem + SM_HBAR), 1);
        *reinterpret_cast<unsigned long long*>(mg_smem + SM_HBAR + 8) = globaltimer_ns() + a.timeout_ns;
        mbar_fence_init();
    }
    // the arguments, once, into shared memory for the phase functions (see mg_args())
    for (int i = tid; i < (int)(sizeof(DlDecodeArgs) / 4); i += MG_THREADS)
        reinterpret_cast<uint32_t*>(mg_smem + SM_ARGS)[i] = reinterpret_cast<const uint32_t*>(&a)[i];
    if (tid < MG_MAX_M) {                                    // running greedy best of this CTA
        float* run = c.misc + MG_BEST_OFF + 4 * MG_MAX_M * 2;
        run[tid * 2] = -INFINITY;
        run[tid * 2 + 1] = __int_as_float(0x7fffffff);
    }
    if (tid < a.M) {
        MgRowInfo ri;
        ri.pos = a.pos[tid];
        ri.slot = a.slot[tid];
        ri.token = a.tokens[tid];
        ri.page = a.block_table[(size_t)ri.slot * a.max_pages + ri.pos / MG_PAGE];
        ri.n_eff = mg_n_eff(ri.pos, a.att_splits);
        ri.pad0 = ri.pad1 = ri.pad2 = 0;
        c.rows[tid] = ri;
    }
    if (tid >= 32 && tid < 48)                               // layer table entry of layer 0 (see the layer loop)
        reinterpret_cast<uint32_t*>(c.lay)[tid - 32] = reinterpret_cast<const uint32_t*>(a.layers)[tid - 32];
    for (int i = tid; i < a.M * (a.head_size >> 1); i += MG_THREADS) {      // RoPE angles (OP_ROPE_LLAMA, nn-cpu-ops.cpp:1090-1120)
        const int m = i / (a.head_size >> 1), pr = i - m * (a.head_size >> 1);
        float sn, cs;
        sincosf((float)a.pos[m] * a.rope_freqs[pr], &sn, &cs);
        c.rope[m * 64 + pr] = make_float2(cs, sn);
    }
    __syncthreads();
    const uint32_t step = ld_volatile_u32(a.step_counter) + 1u;          // nobody writes it before the last grid barrier

    if (warp >= 16 && warp < MG_KV_WARP) { weight_producer(); return; }
    if (warp == MG_KV_WARP) { kv_producer(); return; }

    // ------------------------------------------ consumers ------------------------------------------
    MgSpin sp;
    sp.init(a.timeout_ns, a.debug);
    const int M = a.M, L = a.n_layers, N = a.world;
    const uint32_t ebase = step * (uint32_t)(2 * L);       // LL epochs of this step: ebase + 2l + 1 (wo), ebase + 2l + 2 (w2)
    uint32_t it = 0, ai = 0;
    unsigned nbar = 0;
    int cur = 0;
    // optional timeline: 16 globaltimer stamps per layer (+ 2 at the end) of the CTA selected by trace_cta
    unsigned long long* tr = (a.trace != nullptr && c.cta == a.trace_cta && tid == 0) ? a.trace : nullptr;
#define MG_STAMP(l, k) do { if (tr) { tr[(l) * 32 + (k)] = globaltimer_ns(); if ((k) == 0) c.rows[0].pad0 = (l); } } while (0)
    // do any rows need the cooperative split merge?  (uniform over the grid: same inputs everywhere)
    bool coop = false;
    for (int m = 0; m < M; m++) coop = coop || (c.rows[m].n_eff > MG_SMALL_SPLITS);
    const int att_sel = a.head_size * 8 + (a.n_heads / a.n_kv_heads > 4 ? 2 : 1);
    int R0, R1;

    // The residual stream x is replicated: every CTA rebuilds the row it needs in its norm prologue from the previous x
    // and the ranks' Q80 partial sums (LL words), CTA 0 writes it back for the next prologue (ping-pong buffers; the
    // grid barriers in between order those writes against the reads of the other CTAs).
#pragma unroll 1
    for (int l = 0; l < L; l++) {
        MG_STAMP(l, 0);
        // the consumers take their per-layer pointers from shared memory; the entry of the NEXT layer is fetched now,
        // a whole layer before it is needed (slot (l + 1) & 1 was last read in layer l - 1, grid barriers ago)
        if (tid >= 32 && tid < 48 && l + 1 < L)
            reinterpret_cast<uint32_t*>(c.lay + ((l + 1) & 1))[tid - 32] = reinterpret_cast<const uint32_t*>(a.layers + l + 1)[tid - 32];
        // ---- attention block: norm -> qkv ----
        {
            mg_cta_rows(a.q_dim + 2 * a.kv_dim, c.G, 2, c.cta, R0, R1);
            // layer 0: CTA 0 keeps the embedding rows as x; later layers: the merger CTAs write the merged row for everybody
            float* xo = (l == 0) ? (c.cta == 0 ? a.x[0] : nullptr) : a.x[cur ^ 1];
            if (R1 > R0 || c.cta < MG_MERGERS)
                prologue_norm(MT, l == 0 ? nullptr : a.x[cur], xo, l > 0 ? a.ll_ffn[a.rank] : nullptr,
                              ebase + 2 * l /* w2 of layer l-1 */, c.lay[l & 1].rms_att);
            if (l > 0) cur ^= 1;
        }
        MG_STAMP(l, 1);
        it = gemv_phase<MT>(PH_QKV, l, it, 0u);
        MG_STAMP(l, 2);
        cons_kv_bar();                 // the staging area is free: the KV warp starts on the pages this step does not write
        grid_sync(a.grid_bar, c.G, nbar, sp);
        MG_STAMP(l, 3);
        // ---- attention ----
        cons_kv_bar2();
        switch (att_sel) {
            case 32 * 8 + 1: ai = attention_phase<32, 1>(ai); break;
            case 32 * 8 + 2: ai = attention_phase<32, 2>(ai); break;
            case 64 * 8 + 1: ai = attention_phase<64, 1>(ai); break;
            case 64 * 8 + 2: ai = attention_phase<64, 2>(ai); break;
            case 128 * 8 + 1: ai = attention_phase<128, 1>(ai); break;
            default: ai = attention_phase<128, 2>(ai); break;
        }
        MG_STAMP(l, 4);
        grid_sync(a.grid_bar, c.G, nbar, sp);
        if (coop) {
            attention_merge_coop();
            grid_sync(a.grid_bar, c.G, nbar, sp);
        }
        MG_STAMP(l, 5);
        // ---- wo ----
        mg_cta_rows(a.dim, c.G, 32, c.cta, R0, R1);
        if (R1 > R0) prologue_att(MT, coop);
        MG_STAMP(l, 6);
        it = gemv_phase<MT>(PH_WO, l, it, ebase + 2 * l + 1);
        MG_STAMP(l, 7);
        MG_STAMP(l, 8);
        // ---- feed-forward block: norm -> w1|w3 ----
        {
            mg_cta_rows(2 * a.hid, c.G, 2, c.cta, R0, R1);
            float* xo = a.x[cur ^ 1];
            if (R1 > R0 || c.cta < MG_MERGERS)
                prologue_norm(MT, a.x[cur], xo, a.ll_att[a.rank], ebase + 2 * l + 1, c.lay[l & 1].rms_ffn);
            cur ^= 1;
        }
        MG_STAMP(l, 9);
        it = gemv_phase<MT>(PH_W13, l, it, 0u);
        MG_STAMP(l, 10);
        grid_sync(a.grid_bar, c.G, nbar, sp);
        MG_STAMP(l, 11);
        // ---- w2 ----
        mg_cta_rows(a.dim, c.G, 32, c.cta, R0, R1);
        if (R1 > R0) prologue_hidden(MT);
        MG_STAMP(l, 12);
        it = gemv_phase<MT>(PH_W2, l, it, ebase + 2 * l + 2);
        MG_STAMP(l, 13);
        MG_STAMP(l, 14);
    }

    // ---- final norm -> logits (OP_MATMUL wcls + logits gather, src/llm.cpp:394-432) ----
    mg_cta_rows(a.vocab_loc, c.G, 2, c.cta, R0, R1);
    if (R1 > R0 || c.cta < MG_MERGERS) prologue_norm(MT, a.x[cur], a.x[cur ^ 1], a.ll_ffn[a.rank], ebase + 2 * L, a.rms_final);
    it = gemv_phase<MT>(PH_CLS, 0, it, 0u);
    MG_STAMP(L, 1);
    if (a.greedy) {
        // CTA candidates -> global; CTA 0 reduces them, exchanges the per-rank winners as LL words, picks the token
        if (tid < M) {
            const float* run = c.misc + MG_BEST_OFF + 4 * MG_MAX_M * 2;
            a.cta_best[((size_t)c.cta * MG_MAX_M + tid) * 2] = run[tid * 2];
            a.cta_best[((size_t)c.cta * MG_MAX_M + tid) * 2 + 1] = run[tid * 2 + 1];
        }
    } else if (N > 1) {
        cons_bar();
        if (tid == 0) __threadfence_system();      // this CTA's logits stores to the peers, before the grid barrier
    }
    grid_sync(a.grid_bar, c.G, nbar, sp);
    MG_STAMP(L, 2);
    if (tid == 0) {
        // the last CTA to get here clears the barrier words for the next launch (everybody has left the last spin)
        const unsigned prev = atomicAdd(a.grid_bar + 1, 1u);
        if (prev == (unsigned)c.G - 1u) {
            a.grid_bar[0] = 0u;
            a.grid_bar[1] = 0u;
        }
    }
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
            if (N > 1) {
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
    } else if (N > 1) {
        if (tid < N) st_release_sys(a.flag_logits[tid], step);
    }
    cons_bar();
    if (tid == 0) *a.step_counter = step;
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
    const int mt = a->M == 1 ? 1 : (a->M == 2 ? 2 : 4);         // rows the activation slabs carry
    // prologue_norm parks the merged row (f32) behind its Q80 slabs
    if ((long long)a->M * (a->dim / 32) * MG_XS_STRIDE * 16 + (long long)(a->dim / 128) * mt * XQ_ROW_BYTES > MG_ACT_BYTES) { dl_set_error("decode_mega: M*dim=%lld does not fit the shared-memory row buffer", (long long)a->M * a->dim); return -1; }
    const long long kmax = a->hid > a->dim ? (a->hid > a->q_dim ? a->hid : a->q_dim) : (a->dim > a->q_dim ? a->dim : a->q_dim);
    if (kmax / 128 * mt * XQ_ROW_BYTES > MG_ACT_BYTES) { dl_set_error("decode_mega: activation slabs (%lld B) exceed %d B of shared memory", kmax / 128 * mt * XQ_ROW_BYTES, MG_ACT_BYTES); return -1; }
    if (a->world < 1 || a->world > DL_MAX_PEERS || a->rank < 0 || a->rank >= a->world) { dl_set_error("decode_mega: rank/world"); return -1; }
    if (!a->q80_sync) { dl_set_error("decode_mega: the wo / w2 partial sums always travel as Q80 words (q80_sync must be on)"); return -1; }
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
