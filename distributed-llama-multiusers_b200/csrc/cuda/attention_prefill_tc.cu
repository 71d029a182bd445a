// Causal attention for long prompt chunks on the 5th-gen tensor cores (head size 128): the same job as
// attention_prefill.cu (segments of consecutive positions of one sequence, listed in `meta`), with both products
// as tcgen05.mma and the score / probability / output tiles in TENSOR MEMORY instead of registers.
//
// A CTA owns 128 query rows x one query head and walks the KV history 128 tokens (four pages) at a time:
//   S = Q K^T     SS-mode MMA, M = 128 rows, N = 128 tokens, K = 128: Q staged once (fp16, K-major, 128-byte swizzle);
//                 K pages arrive by TMA tensor copies (cp.async.bulk.tensor.2d over this GPU's KV cache, box 64 x 32,
//                 hardware 128-byte swizzle); S (f32) lands in TMEM, double-buffered
//   softmax       two threads per query row (= TMEM lane), 64 score columns each, read into registers with tcgen05.ld
//                 in one pass (the score buffer is handed back at once); the row maximum is exchanged through shared
//                 memory; exp2 -> fp16 probabilities written back to TMEM with tcgen05.st (every fourth exp2 on the
//                 FMA pipe: the step is MUFU-bound); running maximum with LAZY rescaling: the reference maximum only
//                 moves when a row's maximum grew by more than 2^8, so in the steady state the output accumulator is
//                 never touched by the softmax threads
//   O += P V      TS-mode MMA: A = P straight from TMEM, B = V pages in shared memory as an MN-major operand (V is
//                 [token][head dim] in the cache, i.e. N-contiguous: no transpose anywhere), accumulating in TMEM
// S(j+1) and S(j+2) are issued while the softmax threads are busy with step j, so the tensor core works ahead of them.  Output: O / l -> shared memory -> Q80 cast into the XQ / XH slabs of the wo matmul.
//
// Warp roles (384 threads): warp 0 issues the S MMAs (+ TMEM owner), warp 6 the P V MMAs (an issuing thread spends ~700 clk
// per group of 8 MMAs, so one thread for both would cap a step at ~1500 clk), warps 1-4 and 8-11 softmax (two threads per
// row = TMEM lane, 64 columns each; the row maximum is exchanged through shared memory), warp 5
// K producer and warp 7 V producer (one lane each issues 8 tensor copies of 4 KB per step: 4 pages x 2 halves of the head
// dimension; K and V have their own 3-stage rings, so a K stage is free again as soon as its scores are complete).
// (The first version copied K/V with cp.async in 16-byte pieces into un-swizzled core matrices: 4096 LDGSTS per step
// cost 3500 clk of issue time, three times the MMAs -- scripts/trace_attention_tc.py.)
// TMEM columns: S0 [0,128) S1 [128,256) P0 [256,320) P1 [320,384) O [384,512).
// Reference semantics: multiheadAtt_F32 (src/nn/nn-cpu-ops.cpp:749-784), softmax in f32, output cast as OP_CAST
// (src/llm.cpp:296-301).  KV cache: fp16 pages [page][kv_head][32][128].
#include <cuda.h>
#include "epilogue.cuh"

#define PT_HS 128
#define PT_ROWS 128
#define PT_KV 128                      // tokens per step
#define PT_THREADS 384
#define PT_MAX_SEG 8
#define PT_TILE_BYTES 32768            // 128 x 128 fp16
#define PT_STAGES 3                    // KV tiles in flight (K + V: 64 KB each)
#define PT_OUT_LD 129
#define PT_DEADLINE_NS 4000000000ull

struct PtArgs {
    const float* q;            // [M][n_heads*128]
    const __half* k_cache;     // [pages][n_kv][32][128]
    const __half* v_cache;
    const int* pos;
    const int* slot;
    const int* block_table;
    const int* meta;           // meta[0] = n_seg, then [begin, end) pairs
    int M, n_heads, n_kv_heads, max_pages;
    uint8_t* xq_out;
    uint8_t* xh_out;
    int out_mcap;
    float* y_out;
    long long* stamps;         // debug: clock64 stamps of CTA (0, 0), [tile][8]
    int dbg;                   // timing experiments (results invalid): 1 = no softmax arithmetic, 2 = no KV copies, 4 = no MMAs
    unsigned long long* tr;
};

DL_DEVINL void pt_wait(uint64_t* bar, uint32_t parity, unsigned long long t0) {
    unsigned n = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++n & 0x3FFu) == 0 && globaltimer_ns() - t0 > PT_DEADLINE_NS) __trap();
    }
}
DL_DEVINL void pt_cp_async16(uint32_t smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
DL_DEVINL void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
        "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
DL_DEVINL void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
        "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
        "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
        "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
        "r"(v[31])
        : "memory");
}
// instruction descriptor, kind::f16: A fp16 K-major (or TMEM), B fp16 with selectable major-ness, D f32
DL_DEVINL uint32_t pt_idesc(int M, int N, int b_mn_major) {
    uint32_t d = 0;
    d |= 1u << 4;                                  // D = f32
    d |= (uint32_t)(b_mn_major & 1) << 16;         // b_major: 0 = K, 1 = MN
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;
}
#define PT_STAMP(slot, j) do { if (a.stamps && blockIdx.x == 0 && blockIdx.y == 0 && (j) < 64) a.stamps[(j) * 8 + (slot)] = clock64(); } while (0)

DL_DEVINL float pt_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
DL_DEVINL uint32_t pt_pack_half2(float lo, float hi) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}

// exp2 on the FMA pipe (Cody-Waite split + degree-3 polynomial, relative error 8e-5 -- the result is rounded to fp16,
// 5e-4, right after): the softmax is bound by the 16 MUFU results per clock and SM (128 x 128 exp2 per step = 1024 clk,
// as long as the MMAs of the step), so every fourth element takes this path instead
DL_DEVINL float pt_ex2_poly(float x) {
    const float xc = fmaxf(x, -126.0f);
    const float t = xc + 12582912.0f;              // 1.5 * 2^23: round(x) lands in the low mantissa bits
    const float f = xc - (t - 12582912.0f);        // [-0.5, 0.5]
    const float p = fmaf(fmaf(fmaf(0.0550886838f, f, 0.242604051f), f, 0.693276242f), f, 0.99992894f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

// 32 scores (masked ones are -inf) -> exp2(s * sc - ref) as 16 packed fp16 pairs into TMEM; returns their sum
DL_DEVINL float pt_probs(const uint32_t (&v)[32], float sc, float ref, uint32_t taddr) {
    uint32_t pk[16];
    float sum = 0.0f;
#pragma unroll
    for (int e = 0; e < 32; e += 2) {
        const float pa = pt_ex2(fmaf(__uint_as_float(v[e]), sc, -ref));
        const float xb = fmaf(__uint_as_float(v[e + 1]), sc, -ref);
        const float pb = (e & 2) ? pt_ex2_poly(xb) : pt_ex2(xb);
        sum += pa + pb;
        pk[e >> 1] = pt_pack_half2(pa, pb);
    }
    tmem_st_32x32b_x16(taddr, pk);
    return sum;
}

__global__ void __launch_bounds__(PT_THREADS, 1) attention_prefill_tc_kernel(const __grid_constant__ PtArgs a, const __grid_constant__ CUtensorMap kmap,
                                                                                const __grid_constant__ CUtensorMap vmap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    trace_begin(a.tr);
    const unsigned long long t0 = globaltimer_ns();
    uint8_t* q_s = smem;                                        // Q tile, K-major core matrices
    uint8_t* k_s = smem + PT_TILE_BYTES;                        // [3] K tiles (tokens x head dim, K-major)
    uint8_t* v_s = k_s + PT_STAGES * PT_TILE_BYTES;             // [3] V tiles (tokens x head dim, MN-major for P V)
    float* out_s = reinterpret_cast<float*>(k_s);               // [128][129] f32, reuses the KV stages after the loop
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (1 + 2 * PT_STAGES) * PT_TILE_BYTES);
    uint64_t* k_full = bars;             // [3] transaction bytes of the step's K copies
    uint64_t* k_empty = bars + 3;        // [3] the scores that read the stage are complete
    uint64_t* v_full = bars + 6;         // [3] transaction bytes of the step's V copies
    uint64_t* v_empty = bars + 9;        // [3] P V of the step has retired
    uint64_t* q_full = bars + 12;
    uint64_t* s_full = bars + 13;        // [2] scores complete
    uint64_t* p_full = bars + 15;        // [2] probabilities written (and the score buffer read)
    uint64_t* o_full = bars + 17;        // [2] P V of tile j complete (barrier j & 1)
    uint64_t* s_empty = bars + 19;       // [2] the score buffer has been read into registers
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 21);
    float* mx_s = reinterpret_cast<float*>(bars + 24);          // [2 step parities][2 halves][128 rows] row maxima (and row sums at the end)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int head = blockIdx.y;
    const int kvh = head / (a.n_heads / a.n_kv_heads);
    pdl_launch_dependents();
    pdl_wait();                                                  // q, this chunk's KV rows and meta come from earlier kernels
    // which (segment, 128-row tile) is this CTA?  Large tiles (late rows: long histories) are handed out first.
    int tile = blockIdx.x, begin = 0, end = 0;
    {
        const int n_seg = min(a.meta[0], PT_MAX_SEG);
        int i = 0;
        for (; i < n_seg; i++) {
            begin = a.meta[1 + 2 * i];
            end = a.meta[2 + 2 * i];
            const int nt = (end - begin + PT_ROWS - 1) / PT_ROWS;
            if (tile < nt) { tile = nt - 1 - tile; break; }
            tile -= nt;
        }
        if (i == n_seg) { trace_end(a.tr); return; }
    }
    const int n_rows = end - begin;
    const int r_first = tile * PT_ROWS;
    const int rows_here = min(PT_ROWS, n_rows - r_first);
    const int p0 = a.pos[begin];
    const int* bt = a.block_table + (size_t)a.slot[begin] * a.max_pages;
    const int last_pos = p0 + r_first + rows_here - 1;
    const int seq_end = p0 + n_rows;                             // tokens of this sequence that exist
    const int n_tiles = last_pos / PT_KV + 1;

    if (tid == 0) {
        for (int s = 0; s < PT_STAGES; s++) {
            mbar_init(&k_full[s], 1);
            mbar_init(&k_empty[s], 1);
            mbar_init(&v_full[s], 1);
            mbar_init(&v_empty[s], 1);
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(&s_full[s], 1);
            mbar_init(&s_empty[s], 256);
            mbar_init(&p_full[s], 256);
            mbar_init(&o_full[s], 1);
        }
        mbar_init(q_full, 256);
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc(tmem_holder, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    const uint32_t T_S = tmem_base, T_P = tmem_base + 256, T_O = tmem_base + 384;

    // 128-byte-swizzled operands: rows of 128 B (64 halfs), 8-row groups 1024 B apart; the head dimension is split in
    // two such half tiles 16 KB apart.  K-major (Q, K): LBO unused (1), SBO 1024.  MN-major (V as B of P V): LBO =
    // distance between the two 64-wide halves of N (16 KB), SBO = distance between 8-token groups (1024 B).
    const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    if (warp == 0) {
        // ------------------------------------ S = Q K^T issuer ------------------------------------
        const uint32_t idesc_s = pt_idesc(PT_ROWS, PT_KV, 0);
        auto lo_k = [](const void* p) { return ((smem_u32(p) & 0x3FFFFu) >> 4) | (1u << 16); };
        pt_wait(q_full, 0, t0);
        for (int j = 0; j < n_tiles; j++) {
            pt_wait(&k_full[j % PT_STAGES], (j / PT_STAGES) & 1, t0);
            if (lane == 0) PT_STAMP(0, j);
            if (j >= 2) pt_wait(&s_empty[j & 1], ((j - 2) >> 1) & 1, t0);     // softmax(j-2) has this score buffer in registers
            tc_fence_after();
            if (elect_one()) {
                const uint32_t a_lo = lo_k(q_s), b_lo = lo_k(k_s + (j % PT_STAGES) * PT_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < 8; k++) {      // K = 16 per MMA: 32 B inside the swizzled row, the second half tile after four
                    const uint32_t off = (uint32_t)((k >> 2) * (16384 >> 4) + (k & 3) * 2);
                    uint64_t da, db;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(da) : "r"(a_lo + off), "r"(desc_hi));
                    asm("mov.b64 %0, {%1, %2};" : "=l"(db) : "r"(b_lo + off), "r"(desc_hi));
                    if (!(a.dbg & 4)) umma_f16_ss(T_S + (j & 1) * 128, da, db, idesc_s, k > 0 ? 1u : 0u);
                }
                umma_commit(&s_full[j & 1]);
                umma_commit(&k_empty[j % PT_STAGES]);
            }
            __syncwarp();
            if (lane == 0) PT_STAMP(1, j);
        }
    } else if (warp == 6) {
        // ------------------------------------ O += P V issuer ------------------------------------
        const uint32_t idesc_pv = pt_idesc(PT_ROWS, PT_HS, 1);
        for (int i = 0; i < n_tiles; i++) {
            uint8_t* vz = v_s + (i % PT_STAGES) * PT_TILE_BYTES;
            if ((i + 1) * PT_KV > seq_end) {
                // the step that contains the end of the sequence: rows past it were never written (or not copied at all);
                // their scores are masked, but a zero probability times a non-finite V would poison the product -> zero them
                pt_wait(&v_full[i % PT_STAGES], (i / PT_STAGES) & 1, t0);
                const int r0 = max(0, seq_end - i * PT_KV);
                for (int idx = r0 * 16 + lane; idx < PT_KV * 16; idx += 32) {      // 16 pieces of 16 B per token row (two halves)
                    const int r = idx >> 4, c = idx & 15;
                    *reinterpret_cast<uint4*>(vz + (c >> 3) * 16384 + r * 128 + (c & 7) * 16) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
                __syncwarp();
            }
            pt_wait(&v_full[i % PT_STAGES], (i / PT_STAGES) & 1, t0);
            pt_wait(&p_full[i & 1], (i >> 1) & 1, t0);
            if (lane == 0) PT_STAMP(2, i);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t b_lo = ((smem_u32(vz) & 0x3FFFFu) >> 4) | ((16384u >> 4) << 16);
#pragma unroll
                for (int k = 0; k < 8; k++) {      // 16 tokens per MMA: 8 TMEM columns of P, two 8-token groups (2048 B) of V
                    uint64_t db;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(db) : "r"(b_lo + k * 128), "r"(desc_hi));
                    if (!(a.dbg & 4)) umma_f16_ts(T_O, T_P + (i & 1) * 64 + k * 8, db, idesc_pv, (i > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&o_full[i & 1]);
                umma_commit(&v_empty[i % PT_STAGES]);
            }
            __syncwarp();
        }
    } else if (warp <= 4 || warp >= 8) {
        // ---------------- softmax: two threads per query row (= TMEM lane), 64 score columns each ----------------
        // warps 1-4 take columns [0, 64) of their lane quarter, warps 8-11 columns [64, 128): the MUFU work of a step
        // (128 x 128 exp2) is fixed, but with two warps per scheduler the loads, maxima, packs and stores overlap it
        const int quarter = warp & 3, row = quarter * 32 + lane, hf = warp >= 8 ? 1 : 0;
        const int r_abs = r_first + row;
        {
            const float* qg = a.q + (size_t)(begin + min(r_abs, n_rows - 1)) * a.n_heads * PT_HS + head * PT_HS + hf * 64;
            uint8_t* qd = q_s + hf * 16384 + row * 128;
#pragma unroll 4
            for (int c = 0; c < 8; c++) {       // 16-byte piece c of this half of the row -> swizzled slot c ^ (row & 7)
                const float4 x = reinterpret_cast<const float4*>(qg)[2 * c], y = reinterpret_cast<const float4*>(qg)[2 * c + 1];
                uint4 u;
                u.x = pt_pack_half2(x.x, x.y); u.y = pt_pack_half2(x.z, x.w);
                u.z = pt_pack_half2(y.x, y.y); u.w = pt_pack_half2(y.z, y.w);
                *reinterpret_cast<uint4*>(qd + ((c ^ (row & 7)) << 4)) = u;
            }
            fence_proxy_async_smem();
            mbar_arrive(q_full);
        }
        const int pos_r = p0 + r_abs;
        const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
        const float sc = 1.4426950408889634f * rsqrtf((float)PT_HS);      // scores in the log2 domain
        float m_used = -INFINITY, l = 0.0f;
        uint32_t v0[32], v1[32];
        for (int j = 0; j < n_tiles; j++) {
            pt_wait(&s_full[j & 1], (j >> 1) & 1, t0);
            if (tid == 32) PT_STAMP(3, j);
            tc_fence_after();
            const uint32_t t_s = T_S + lane_base + (j & 1) * 128 + hf * 64;
            const int tok0 = j * PT_KV + hf * 64;
            if (a.dbg & 1) { tc_fence_before(); mbar_arrive(&s_empty[j & 1]); mbar_arrive(&p_full[j & 1]); continue; }
            tmem_ld_32x32b_x32(t_s, v0);
            tmem_ld_32x32b_x32(t_s + 32, v1);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&s_empty[j & 1]);                    // the scores of step j+2 may overwrite the buffer now
            const int lim = pos_r - tok0;                    // columns e <= lim of this half are visible to this row
            float ma = -INFINITY, mb = -INFINITY;
            if (lim >= 63) {
#pragma unroll
                for (int e = 0; e < 32; e += 2) {
                    ma = fmaxf(ma, fmaxf(__uint_as_float(v0[e]), __uint_as_float(v1[e])));
                    mb = fmaxf(mb, fmaxf(__uint_as_float(v0[e + 1]), __uint_as_float(v1[e + 1])));
                }
            } else {
#pragma unroll
                for (int e = 0; e < 32; e++) {
                    if (e > lim) v0[e] = 0xff800000u;        // -inf
                    if (e + 32 > lim) v1[e] = 0xff800000u;
                    ma = fmaxf(ma, fmaxf(__uint_as_float(v0[e]), __uint_as_float(v1[e])));
                }
            }
            // row maximum over both halves: exchange through shared memory (slot of step parity: the next write to it
            // comes two steps later, after everybody has passed the barrier of the step in between)
            float mx = fmaxf(ma, mb);
            mx_s[((j & 1) * 2 + hf) * PT_ROWS + row] = mx;
            asm volatile("bar.sync 2, 256;" ::: "memory");
            mx = fmaxf(mx, mx_s[((j & 1) * 2 + (hf ^ 1)) * PT_ROWS + row]);
            if (tid == 32) PT_STAMP(4, j);
            mx *= sc;
            const bool need = mx > m_used + 8.0f;
            if (__any_sync(0xffffffffu, need)) {
                const float m_new = need ? mx : m_used;
                const float alpha = need ? pt_ex2(m_used - m_new) : 1.0f;
                if (j > 0) {      // the accumulated output moves to the new reference (rare after the first tiles)
                    pt_wait(&o_full[(j - 1) & 1], ((j - 1) >> 1) & 1, t0);
                    tc_fence_after();
                    uint32_t w[32];
#pragma unroll 1
                    for (int ch = 0; ch < 2; ch++) {
                        tmem_ld_32x32b_x32(T_O + lane_base + hf * 64 + ch * 32, w);
                        tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 32; e++) w[e] = __float_as_uint(__uint_as_float(w[e]) * alpha);
                        tmem_st_32x32b_x32(T_O + lane_base + hf * 64 + ch * 32, w);
                    }
                    tmem_st_wait();
                }
                l *= alpha;
                m_used = m_new;
            }
            const float ref = (m_used == -INFINITY) ? 0.0f : m_used;
            const uint32_t t_p = T_P + lane_base + (j & 1) * 64 + hf * 32;
            if (j >= 2) {      // P V(j-2) has consumed this probability buffer (long done; the wait keeps it a guarantee)
                pt_wait(&o_full[j & 1], ((j - 2) >> 1) & 1, t0);
                tc_fence_after();
            }
            l += pt_probs(v0, sc, ref, t_p);
            l += pt_probs(v1, sc, ref, t_p + 16);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&p_full[j & 1]);
            if (tid == 32) PT_STAMP(5, j);
        }
        // ---- finish: O / l -> shared memory -> Q80 cast into the slabs of the wo matmul ----
        pt_wait(&o_full[(n_tiles - 1) & 1], ((n_tiles - 1) >> 1) & 1, t0);
        tc_fence_after();
        // the two halves of the row sum, through the exchange slots the LAST step did not use (their previous readers
        // have all passed that step's barrier)
        const int fs = (n_tiles & 1) * 2;
        mx_s[(fs + hf) * PT_ROWS + row] = l;
        asm volatile("bar.sync 2, 256;" ::: "memory");
        l += mx_s[(fs + (hf ^ 1)) * PT_ROWS + row];
        const float inv = l > 0.0f ? 1.0f / l : 0.0f;
#pragma unroll 1
        for (int ch = 0; ch < 2; ch++) {
            tmem_ld_32x32b_x32(T_O + lane_base + hf * 64 + ch * 32, v0);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; e++) out_s[row * PT_OUT_LD + hf * 64 + ch * 32 + e] = __uint_as_float(v0[e]) * inv;
        }
        __syncwarp();                                            // a warp only re-reads the 32 rows x 64 columns it wrote
        for (int rr = 0; rr < 32; rr++) {
            const int rc = r_first + quarter * 32 + rr;
            if (rc >= n_rows) break;
            const int m = begin + rc;
#pragma unroll
            for (int b = 0; b < 2; b++) {
                const float val = out_s[(quarter * 32 + rr) * PT_OUT_LD + hf * 64 + b * 32 + lane];
                q80_block_store(val, lane, m, (head * PT_HS) / 32 + hf * 2 + b, a.out_mcap, a.xq_out, a.xh_out);
                if (a.y_out != nullptr) a.y_out[(size_t)m * a.n_heads * PT_HS + head * PT_HS + hf * 64 + b * 32 + lane] = val;
            }
        }
    } else if (warp == 5 || warp == 7) {
        // ------------------------------------ K producer (warp 5) / V producer (warp 7), one lane each ------------------------------------
        if (lane == 0) {
            const bool is_k = warp == 5;
            const uint64_t map = is_k ? reinterpret_cast<uint64_t>(&kmap) : reinterpret_cast<uint64_t>(&vmap);
            uint64_t* full = is_k ? k_full : v_full;
            uint64_t* empty = is_k ? k_empty : v_empty;
            uint8_t* ring = is_k ? k_s : v_s;
            const int total_pages = (seq_end + 31) / 32;
            // the page ids of a step are fetched one step ahead, so the table reads (dependent global loads) overlap the
            // wait for a free stage instead of delaying the copies
            int pg[4], pg_next[4];
#pragma unroll
            for (int p4 = 0; p4 < 4; p4++) pg_next[p4] = (p4 < total_pages) ? bt[p4] : 0;
            for (int j = 0; j < n_tiles; j++) {
                const int st = j % PT_STAGES;
#pragma unroll
                for (int p4 = 0; p4 < 4; p4++) {
                    pg[p4] = pg_next[p4];
                    pg_next[p4] = (4 * (j + 1) + p4 < total_pages) ? bt[4 * (j + 1) + p4] : 0;
                }
                if (j >= PT_STAGES) pt_wait(&empty[st], ((j / PT_STAGES) & 1) ^ 1, t0);
                if (is_k) PT_STAMP(6, j);
                const int n_pg = (a.dbg & 2) ? 0 : min(4, (seq_end - j * PT_KV + 31) / 32);      // pages of this step that exist
                mbar_arrive_expect_tx(&full[st], (uint32_t)n_pg * 2u * 4096u);
                const uint32_t dst = smem_u32(ring + st * PT_TILE_BYTES), bar = smem_u32(&full[st]);
#pragma unroll
                for (int p4 = 0; p4 < 4; p4++) {
                    if (p4 >= n_pg) break;
                    const int y = (pg[p4] * a.n_kv_heads + kvh) * 32;      // first row of the page in the [rows][128] view
#pragma unroll
                    for (int h = 0; h < 2; h++)
                        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                                         dst + h * 16384 + p4 * 4096),
                                     "l"(map), "r"(h * 64), "r"(y), "r"(bar)
                                     : "memory");
                }
                if (is_k) PT_STAMP(7, j);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    trace_end(a.tr);
}

extern void dl_set_error(const char* fmt, ...);
static int g_pt_dbg = 0;
static long long* g_pt_stamps = nullptr;
extern "C" void dl_debug_set_pt_mode(int m) { g_pt_dbg = m; }
extern "C" void dl_debug_set_pt_trace(long long* p) { g_pt_stamps = p; }

typedef CUresult (*PtEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// [rows][128] fp16 view of a KV cache on THIS GPU (rows = pages x kv heads x 32), box = 64 columns x 32 rows (one page of
// one head, half of the head dimension), 128-byte swizzle: the image tcgen05.mma reads as a swizzled K- or MN-major tile
static int pt_make_map(CUtensorMap* m, const void* base, long long rows) {
    static PtEncodeTiled enc = nullptr;
    if (enc == nullptr) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || fn == nullptr) {
            dl_set_error("attention_prefill_tc: cuTensorMapEncodeTiled not available");
            return -1;
        }
        enc = (PtEncodeTiled)fn;
    }
    const cuuint64_t dims[2] = {PT_HS, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {PT_HS * 2};
    const cuuint32_t box[2] = {64, 32}, es[2] = {1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { dl_set_error("attention_prefill_tc: cuTensorMapEncodeTiled failed (%d)", (int)r); return -1; }
    return 0;
}

extern "C" int dl_attention_prefill_tc(const float* q, int M, int n_heads, int n_kv_heads, int head_size, const void* k_cache,
                                       const void* v_cache, const int* pos, const int* slot, const int* block_table,
                                       int max_pages, int page_size, const int* meta, uint8_t* xq_out, uint8_t* xh_out,
                                       int out_mcap, float* y_out, int n_cache_pages, int pdl, void* stream) {
    if (page_size != 32 || head_size != PT_HS) { dl_set_error("dl_attention_prefill_tc: page size 32 and head size 128 only"); return -1; }
    if (n_heads % n_kv_heads != 0) { dl_set_error("dl_attention_prefill_tc: bad head grouping"); return -1; }
    if (n_cache_pages < 1) { dl_set_error("dl_attention_prefill_tc: n_cache_pages < 1"); return -1; }
    PtArgs a;
    a.q = q; a.k_cache = (const __half*)k_cache; a.v_cache = (const __half*)v_cache; a.pos = pos; a.slot = slot;
    a.block_table = block_table; a.meta = meta; a.M = M; a.n_heads = n_heads; a.n_kv_heads = n_kv_heads;
    a.max_pages = max_pages; a.xq_out = xq_out; a.xh_out = xh_out; a.out_mcap = out_mcap; a.y_out = y_out;
    a.dbg = g_pt_dbg;
    a.stamps = g_pt_stamps;
    a.tr = dl_trace_next(TR_ATT);
    alignas(64) CUtensorMap kmap, vmap;
    const long long rows = (long long)n_cache_pages * n_kv_heads * 32;
    if (pt_make_map(&kmap, k_cache, rows) != 0 || pt_make_map(&vmap, v_cache, rows) != 0) return -1;
    const size_t smem = (1 + 2 * PT_STAGES) * PT_TILE_BYTES + 24 * 8 + 4 * PT_ROWS * 4;
    static bool set = false;
    if (!set) {
        cudaError_t e = cudaFuncSetAttribute(attention_prefill_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { dl_set_error("attention_prefill_tc smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((M + PT_ROWS - 1) / PT_ROWS + PT_MAX_SEG), (unsigned)n_heads);      // upper bound of the tiles of all segments
    cfg.blockDim = dim3(PT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    if (pdl) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, attention_prefill_tc_kernel, a, kmap, vmap);
    if (e != cudaSuccess) { dl_set_error("attention_prefill_tc launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}
