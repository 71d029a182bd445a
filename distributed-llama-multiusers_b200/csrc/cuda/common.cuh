// Shared device helpers for the sm_100a kernels: mbarrier, bulk async copy (TMA 1-D), PDL,
// cluster / DSMEM, tcgen05 + TMEM wrappers, warp reductions, system-scope flag primitives.
// Everything is inline PTX; no CUTLASS dependency.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cooperative_groups.h>
#include <stdint.h>

namespace cg = cooperative_groups;

#define DL_DEVINL __device__ __forceinline__

// --------------------------------------------------------------------------------------------
// Device data formats (see docs/FORMATS.md)
// --------------------------------------------------------------------------------------------
// Q40T: a [d][n] Q40 matrix re-tiled for bulk copies.  Tiles of 128 rows x 128 K, ordered
// [row_tile][k_tile]; each tile is 9216 contiguous bytes:
//   [0,8192)    nibble bytes: row r, block kb (4 blocks of 32 per tile row) -> 16 bytes at
//               r*64 + (((kb + (r>>1)) & 3) << 4)      (rotation makes 16-byte smem reads
//               conflict-free both for lane==row and lane==(row%8,kb) mappings)
//   [8192,9216) fp16 scales [r][kb]
// Nibble order inside a block is the file's: byte j = elem j (low) | elem j+16 (high) << 4.
#define Q40T_TILE_ROWS 128
#define Q40T_TILE_K 128
#define Q40T_NIB_BYTES 8192
#define Q40T_SCALE_OFF 8192
#define Q40T_TILE_BYTES 9216

// XQ: Q80 activations for the dp4a GEMV, per k-tile slabs [kt][Mcap][160]:
//   128 int8 (4 blocks x 32, natural order) | 4 x f32 scale (fp16-rounded) | 4 x int32 (8*sum(q))
#define XQ_ROW_BYTES 160

// XH: fp16 activations for the tcgen05 GEMM in UMMA no-swizzle K-major core-matrix layout, per
// k-tile slabs [kt][Ncap/8][16 chunks][8 rows][8 halves]; K order inside a 32-block is the
// dequantizer's natural unpack order (see xh_slot()).
#define XH_ROWGROUP_BYTES 2048   // 8 rows x 128 K x 2 B

DL_DEVINL uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// element e (0..31) of a Q40/Q80 block -> (chunk w in 0..3, slot s in 0..7) of the XH / A-tile
// K order.  Word w of the 16 nibble bytes unpacks to
//   [e(4w), e(4w+2), e(4w+1), e(4w+3), e(4w+16), e(4w+18), e(4w+17), e(4w+19)].
DL_DEVINL void xh_slot(int e, int& w, int& s) {
    int hi = e >> 4;
    int e2 = e & 15;
    w = e2 >> 2;
    int r = e2 & 3;
    int rr = (r == 1) ? 2 : ((r == 2) ? 1 : r);
    s = hi * 4 + rr;
}

// --------------------------------------------------------------------------------------------
// optional per-launch timeline (debug): slot[0] = min start, slot[1] = max end (globaltimer ns)
// --------------------------------------------------------------------------------------------
DL_DEVINL unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
DL_DEVINL void trace_begin(unsigned long long* tr) {
    if (tr != nullptr && threadIdx.x == 0) atomicMin(tr, globaltimer_ns());
}
DL_DEVINL void trace_end(unsigned long long* tr) {
    if (tr != nullptr && threadIdx.x == 0) atomicMax(tr + 1, globaltimer_ns());
}
// host side (prep.cu): next trace slot for a launch of `kind`, or NULL when tracing is off
unsigned long long* dl_trace_next(int kind);
enum { TR_EMBED = 0, TR_RNQ = 1, TR_GEMV = 2, TR_ATT = 3, TR_ARGMAX = 4, TR_TC = 5, TR_OTHER = 6 };

// --------------------------------------------------------------------------------------------
// warp reductions
// --------------------------------------------------------------------------------------------
DL_DEVINL float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
DL_DEVINL float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
DL_DEVINL int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// --------------------------------------------------------------------------------------------
// mbarrier
// --------------------------------------------------------------------------------------------
DL_DEVINL void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
DL_DEVINL void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
DL_DEVINL void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
DL_DEVINL void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
DL_DEVINL bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// non-blocking probe (test_wait never suspends the thread): lets several lanes of ONE warp poll different
// barriers without the first unready one stalling the rest
DL_DEVINL bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
DL_DEVINL void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// wait with a sleep between probes: for roles that share an SM sub-partition with busy math warps, so
// their polling does not eat the issue slots (costs up to `ns` of wake-up latency)
DL_DEVINL void mbar_wait_sleep(uint64_t* bar, uint32_t parity, uint32_t ns) {
    while (!mbar_try_wait(bar, parity)) {
        if (ns) __nanosleep(ns);
    }
}

// ---- cluster-scope variants (DSMEM hand-off between the CTAs of a thread-block cluster) ----
DL_DEVINL uint32_t mapa_shared(uint32_t cta_smem_addr, uint32_t cta_rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_smem_addr), "r"(cta_rank));
    return r;
}
// arrive (release, cluster scope) on an mbarrier that lives in another CTA of the cluster
DL_DEVINL void mbar_arrive_remote(uint32_t cluster_smem_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_smem_addr) : "memory");
}
DL_DEVINL void st_shared_cluster_f32(uint32_t cluster_smem_addr, float v) {
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(cluster_smem_addr), "f"(v) : "memory");
}
DL_DEVINL void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}
DL_DEVINL void fence_acq_rel_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }

// generic-proxy smem writes -> visible to the async proxy (bulk copy engine, tcgen05.mma)
DL_DEVINL void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// --------------------------------------------------------------------------------------------
// bulk async copy global -> shared (1-D TMA; SASS: UBLKCP), completion on an mbarrier
// --------------------------------------------------------------------------------------------
DL_DEVINL void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// same, with an L2 evict-first policy for once-per-token weight streams
DL_DEVINL uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
DL_DEVINL void bulk_g2s_hint(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
        : "memory");
}

// --------------------------------------------------------------------------------------------
// programmatic dependent launch
// --------------------------------------------------------------------------------------------
DL_DEVINL void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
DL_DEVINL void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// --------------------------------------------------------------------------------------------
// system-scope flags (peer-memory signalling)
// --------------------------------------------------------------------------------------------
DL_DEVINL void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DL_DEVINL uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
DL_DEVINL uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
DL_DEVINL uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// wait until *p (written by a peer with st.release.sys) has reached `target` (wrap-safe)
DL_DEVINL void wait_flag_ge(const uint32_t* p, uint32_t target) {
    while ((int32_t)(ld_acquire_sys(p) - target) < 0) {
    }
}

// --------------------------------------------------------------------------------------------
// LL ("low latency") exchange words: {f32 value bits, epoch} in one aligned 8-byte store.  Such a store is
// atomic, so a reader that finds the expected epoch in the upper half also has the value: no fence after
// the peer stores, no counter, no separate flag -- the cost of a hand-off is one NVLink store latency.
// (Same idea as NCCL's LL protocol.)  Epochs increase monotonically, so a stale word can never match.
// --------------------------------------------------------------------------------------------
DL_DEVINL void ll_store(float* base, size_t elem, float v, uint32_t epoch) {
    reinterpret_cast<uint2*>(base)[elem] = make_uint2(__float_as_uint(v), epoch);
}
DL_DEVINL uint4 ld_volatile_v4(const void* p) {
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
// four consecutive values starting at element `elem` (multiple of 4); spins until all four carry `epoch`
DL_DEVINL float4 ll_load4(const float* base, size_t elem, uint32_t epoch) {
    const uint2* p = reinterpret_cast<const uint2*>(base) + elem;
    uint4 a, b;
    while (true) {
        a = ld_volatile_v4(p);
        b = ld_volatile_v4(p + 2);
        if (a.y == epoch && a.w == epoch && b.y == epoch && b.w == epoch) break;
        __nanosleep(100);      // thousands of threads re-reading L2 would slow down the very stores they wait for
    }
    return make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
}
// partial-sum read shared by the consumers: plain f32 or LL words
DL_DEVINL float4 part_load4(const float* base, size_t elem, int ll, uint32_t epoch) {
    return ll ? ll_load4(base, elem, epoch) : *reinterpret_cast<const float4*>(base + elem);
}

// --------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// --------------------------------------------------------------------------------------------
DL_DEVINL void tmem_alloc(uint32_t* smem_holder, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_holder)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
DL_DEVINL void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// one lane of a converged warp (elect.sync)
DL_DEVINL bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}
DL_DEVINL void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DL_DEVINL void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (fp16/bf16 inputs, f32 accumulate)
DL_DEVINL void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]: A lives in tensor memory (lane = row, two fp16 K elements per
// 32-bit column), so only the (small) B operand crosses the shared-memory crossbar
DL_DEVINL void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 16 registers per thread -> 32 lanes x 16 consecutive 32-bit columns of tensor memory (STTM)
DL_DEVINL void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
DL_DEVINL void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// all previously issued MMAs of this thread complete -> arrive(1) on the mbarrier
DL_DEVINL void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// 32 lanes x 32-bit x 8 consecutive columns -> 8 registers per thread
DL_DEVINL void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr)
                 : "memory");
}
DL_DEVINL void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor, K-major, SWIZZLE_NONE ("interleave"): core matrix = 8 rows x
// 16 B stored as 128 contiguous bytes; LBO = byte distance between the two K core matrices of
// one K=16 MMA, SBO = byte distance between consecutive 8-row groups.  (Bit layout as in
// cute/arch/mma_sm100_desc.hpp: addr>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1
// [46,48), layout_type [61,64) = 0.)
DL_DEVINL uint64_t umma_desc_kmajor_noswizzle(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// instruction descriptor for kind::f16, A and B fp16 K-major, D f32
DL_DEVINL uint32_t umma_idesc_f16_f32(int M, int N) {
    uint32_t d = 0;
    d |= 1u << 4;                       // c_format = F32
    d |= 0u << 7;                       // a_format = F16
    d |= 0u << 10;                      // b_format = F16
    d |= 0u << 15;                      // a_major = K
    d |= 0u << 16;                      // b_major = K
    d |= (uint32_t)(N >> 3) << 17;      // n_dim
    d |= (uint32_t)(M >> 4) << 24;      // m_dim
    return d;
}
