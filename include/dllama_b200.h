/* b200-dllama C ABI: every CUDA entry point of libdllama_b200_cuda.so.
 *
 * All pointers are device pointers unless stated otherwise; `stream` is a cudaStream_t.  Functions
 * return 0 on success or a negative/CUDA error code; dl_last_error() gives the message.
 * Citations name the reference component each entry point replaces (paths relative to the
 * reference tree).
 */
#ifndef DLLAMA_B200_H
#define DLLAMA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DL_MAX_PEERS 8

/* ---- epilogue selector for the Q40 GEMV / GEMM kernels ---------------------------------- */
enum {
    DL_EPI_STORE = 0,     /* f32 store (optionally Q80 round-trip) to 1..8 (peer) buffers + flags  */
    DL_EPI_QKV_ROPE = 1,  /* RoPE on q,k rows; q -> buffer; k,v -> paged KV cache                   */
    DL_EPI_SILU_MUL = 2,  /* rows interleaved (gate,up): act(g)*u -> Q80 activation slabs           */
};

typedef struct {
    int kind;
    /* DL_EPI_STORE: dst[p][dst_off + m*ld + row] for p < n_dst  (reference: OP_CAST y->ZQ slice +
     * SYNC_NODE_SLICES, src/llm.cpp:308-314,381-387; logits gather :426-432) */
    float* dst[DL_MAX_PEERS];
    int n_dst;
    int ld;
    long long dst_off;
    int q80_roundtrip;
    const float* residual;            /* NULL, or value += residual[m*residual_ld + row] (TP=1 residual add) */
    int residual_ld;
    uint32_t* flag[DL_MAX_PEERS];     /* flag word to release on each destination (NULL: none) */
    const uint32_t* step_counter;     /* epoch = *step_counter * flag_mul + flag_add           */
    uint32_t flag_mul, flag_add;
    uint32_t* cta_counter;            /* zero-initialised, self-resetting completion counter   */
    /* DL_EPI_QKV_ROPE (reference: OP_ROPE_LLAMA + OP_SHIFT, src/llm.cpp:259-286) */
    float* q_out;                     /* [M][q_dim] */
    int q_dim, kv_dim, head_size;
    const float* rope_freqs;          /* [head_size/2] */
    const int* pos;                   /* [M] */
    const int* slot;                  /* [M] */
    const int* block_table;           /* [n_slots][max_pages] */
    int max_pages, page_size, n_kv_heads;
    void* k_cache;                    /* [n_pages][n_kv_heads][page_size][head_size] */
    void* v_cache;
    int kv_f32;                       /* 1: f32 cache, 0: fp16 cache */
    /* DL_EPI_SILU_MUL (reference: OP_SILU + OP_MUL + OP_CAST d->dq, src/llm.cpp:355-374) */
    uint8_t* xq_out;                  /* XQ slabs of the hidden activation (may be NULL) */
    uint8_t* xh_out;                  /* XH slabs (may be NULL) */
    int out_mcap;                     /* row capacity of the XQ/XH slabs */
    int act_gelu;                     /* 0 = SiLU, 1 = tanh-GELU */
    /* DL_EPI_STORE, tensor-parallel partial sums: 1 = dst[] are arrays of {f32 bits, epoch} 8-byte words (LL
     * protocol: no fence, no flags; step_counter / flag_mul / flag_add give the epoch), 0 = plain f32 */
    int ll;
} DlEpilogue;

/* Fused activation producer for dl_gemv_q40: the kernel itself computes
 *   x = x_in + sum_src partial[src];  y = rmsnorm(x) * norm_w;  Q80(y)
 * for its own K range straight into shared memory (no global round trip, no separate kernel), after
 * waiting on the peers' release flags.  The CTAs of row tile 0 also write x to x_out.
 * (reference: OP_MERGE_ADD + OP_INV_RMS + OP_RMS_NORM + OP_CAST, src/llm.cpp:204-239,317-342) */
typedef struct {
    const float* x_in;                /* [M][dim] residual stream */
    float* x_out;                     /* [M][dim] or NULL */
    const float* partials;            /* [n_src][..] partial sums, src_stride floats apart; may be NULL */
    int n_src;
    long long src_stride;
    const uint32_t* flags;            /* [n_src] release flags to wait on, or NULL */
    const uint32_t* step_counter;
    uint32_t flag_mul, flag_add;
    const float* norm_w;              /* [dim] */
    float eps;
    int ll;                           /* 1: partials are LL words {f32 bits, epoch}; wait on the words, not on flags */
} DlNormProlog;

const char* dl_last_error(void);
int dl_device_sm_count(void);

/* ---- weight preparation ------------------------------------------------------------------ */
/* bytes of a Q40T image of a [d][n] matrix (d padded to 128) */
long long dl_q40t_bytes(int d, int n);
/* scatter `rows` raw Q40 rows (file layout, n/32 blocks of 18 B each) into a Q40T image at
 * destination rows dst_row0 + i*dst_row_step.  (reference: splitRowMatmulWeight /
 * splitColMatmulWeight + NnCpuDevice weight copy, src/nn/nn-core.cpp:270-303) */
int dl_q40_repack(const uint8_t* raw, int rows, int n, uint8_t* q40t, int d_pad, int dst_row0,
                  int dst_row_step, void* stream);
/* Q40T -> f32 [d][n] (tests, cuBLAS baseline path) */
int dl_q40t_dequant(const uint8_t* q40t, int d, int n, float* out, void* stream);

/* ---- activation producers ---------------------------------------------------------------- */
/* x_out = x_in + sum_src partial[src] ; y = rmsnorm(x_out)*w ; y -> Q80 -> XQ and/or XH slabs.
 * Waits (acquire.sys) on flags[src] >= epoch before reading partials when flags != NULL.
 * (reference: OP_MERGE_ADD + OP_INV_RMS + OP_RMS_NORM + OP_CAST, src/llm.cpp:204-239,317-342) */
int dl_reduce_norm_quant(const float* x_in, float* x_out, const float* partials /*f32, or LL words when ll*/, int n_src, int src_stride,
                         const uint32_t* flags, const uint32_t* step_counter, uint32_t flag_mul,
                         uint32_t flag_add, int ll, const float* norm_w, float eps, int dim, int M, int mcap,
                         uint8_t* xq_out, uint8_t* xh_out, float* y_out, int pdl, void* stream);
/* plain f32 [M][K] -> XQ / XH (no norm); (reference: OP_CAST F32->Q80, nn-cpu-ops.cpp:1221-1235) */
int dl_quant_q80(const float* x, int M, int K, int mcap, uint8_t* xq_out, uint8_t* xh_out, void* stream);
/* embedding gather: x[m][:] = table[token[m]][:]  (reference: OP_EMBEDDING, nn-cpu-ops.cpp:880-892) */
int dl_embedding(const float* table, const int* tokens, float* x, int M, int dim, uint32_t* step_counter,
                 void* stream);

/* per-row sampling on the device (reference: Sampler::sample, src/tokenizer.cpp:382-510): temperature 0 -> first
 * argmax; else softmax(logits / temperature) and top-p nucleus (0 < topp < 1) or multinomial, drawn with coin[m] in
 * [0,1) (the caller's per-request RNG).  All arrays are device pointers with M entries; one CTA per row. */
int dl_sample_rows(const float* logits, int M, int vocab, int ld, const float* temperature, const float* topp,
                   const float* coin, int* tokens_out, void* stream);

/* F32 x F32 -> F32 matmul (f32-weights / f32-buffer path; reference: matmul_F32_F32_F32, nn-cpu-ops.cpp:182-220):
 * y[m][r] (+)= sum_k w[r*ldw + k] * x[m*ldx + k], plain f32 FMAs; n, ldw, ldx multiples of 4, w and x 16-byte aligned */
int dl_gemm_f32(const float* w, int d, int n, long long ldw, const float* x, int M, int ldx, float* y, int ldy,
                int accumulate, void* stream);

/* The other kernels of the f32 path (no Q80 cast anywhere; with dl_gemm_f32, dl_reduce_norm_quant (y_out), dl_attention,
 * dl_embedding and dl_argmax they cover the whole forward pass):
 * RoPE on the q and k rows of qkv[M][q_dim + 2 kv_dim], q -> q_out, k / v -> the paged KV cache
 * (reference: OP_ROPE_LLAMA + OP_SHIFT, nn-cpu-ops.cpp:1090-1120,1253-1275) */
int dl_f32_rope_kv(const float* qkv, int M, int q_dim, int kv_dim, int head_size, int n_kv_heads, const float* freqs,
                   const int* pos, const int* slot, const int* block_table, int max_pages, float* q_out,
                   void* k_cache, void* v_cache, int kv_f32, void* stream);
/* out = act(g) * u, act = SiLU or tanh-GELU (reference: OP_SILU / OP_GELU + OP_MUL, nn-cpu-ops.cpp:1058-1078,1163-1183) */
int dl_f32_silu_mul(const float* g, const float* u, float* out, long long n, int gelu, void* stream);
/* store src[M][w] into every destination buffer (peer memory) at dst_off + m*ld_dst and release one flag per destination
 * with epoch *step_counter*flag_mul+flag_add (reference: SYNC_NODE_SLICES of f32 slices, nn-network.cpp:537-569; the
 * merge-add of nn-cpu-ops.cpp:835-852 is done by the consumer, dl_reduce_norm_quant) */
int dl_f32_push(const float* src, int M, int w, int ld_src, float* const* dst /*host array*/, int n_dst, long long dst_off,
                int ld_dst, uint32_t* const* flags /*host array or NULL*/, const uint32_t* step_counter, uint32_t flag_mul,
                uint32_t flag_add, uint32_t* cta_counter /*zero-initialised, self-resetting*/, void* stream);

/* ---- Q40 x Q80 matmuls ------------------------------------------------------------------- */
/* dp4a GEMV, M <= 8 rows: bulk-copy fed, cluster split-K (reference: matmul_Q80_Q40_F32,
 * src/nn/nn-cpu-ops.cpp:222-440) */
int dl_gemv_q40(const uint8_t* w_q40t, int d, int n, const uint8_t* xq /*XQ slabs with mcap == M; unused when norm != NULL*/,
                int M, int row_split /*1,2,4*/, int k_split /*1..8, cluster size*/, int tiles_per_cta /*>=1, only with k_split==1*/,
                const DlNormProlog* norm /*host, NULL = read xq*/,
                const DlEpilogue* epi /*host*/, int pdl, void* stream);
/* shared-memory bytes the GEMV needs for (n, M, k_split); callers pick k_split so that it fits */
long long dl_gemv_q40_smem_bytes(int n, int M, int k_split);
/* tcgen05 GEMM, M up to 256 per tile: Q40 dequantised to fp16 in the smem pipeline, UMMA with
 * TMEM accumulators (reference: llamafile_sgemm Q40xQ80, src/nn/llamafile/sgemm.cpp:593-783) */
int dl_gemm_q40_tc(const uint8_t* w_q40t, int d, int n, const uint8_t* xh, int ncap, int M, int k_split,
                   const DlEpilogue* epi /*host*/, int pdl, void* stream);
/* prompt-sized batches, two phases: the Q40T image is expanded once into WH (fp16 tiles of 128 x 128 in the UMMA
 * K-major core-matrix image, dl_wh_bytes(d, n) bytes), then a persistent SS-mode tcgen05 GEMM with CTA pairs
 * (cta_group::2, 256 x 256 tiles, double-buffered TMEM accumulators) runs on it; same epilogues as above.
 * force_cg: 0 = CTA pairs when the number of 128-row tiles is even, 1 = single CTAs */
long long dl_wh_bytes(int d, int n);
int dl_q40t_to_wh(const uint8_t* w_q40t, int d, int n, uint8_t* wh, void* stream);
int dl_gemm_wh_tc(const uint8_t* wh, int d, int n, const uint8_t* xh, int ncap, int M, const DlEpilogue* epi /*host*/,
                  int force_cg, int pdl, void* stream);

/* ---- attention ---------------------------------------------------------------------------- */
/* per-row paged GQA attention over positions 0..pos[m] (decode rows and prefill-chunk rows alike);
 * output Q80-quantised into XQ/XH slabs for the wo matmul and/or f32.
 * scratch: >= dl_attention_scratch_bytes(); (reference: OP_MULTIHEAD_ATT, nn-cpu-ops.cpp:749-784,1135-1161) */
long long dl_attention_scratch_bytes(int M, int n_heads, int head_size, int max_splits);
int dl_attention(const float* q, int M, int n_heads, int n_kv_heads, int head_size, const void* k_cache,
                 const void* v_cache, int kv_f32, const int* pos, const int* slot, const int* block_table,
                 int max_pages, int page_size, int n_splits, void* scratch, uint8_t* xq_out, uint8_t* xh_out,
                 int out_mcap, float* y_out, const int* row_skip /*device, NULL or [M]: non-zero rows are left to
                 dl_attention_prefill*/, int pdl, void* stream);

/* causal attention of prompt chunks on the tensor cores (mma.sync fp16, f32 accumulate).  meta (device):
 * meta[0] = n_seg (<= 8), then [begin, end) row ranges; each range is consecutive positions of ONE sequence
 * (slot[begin], first position pos[begin]).  Rows outside the ranges are not touched (dl_attention with row_skip
 * serves them).  fp16 KV cache, head_size 64 or 128.  Same outputs as dl_attention. */
int dl_attention_prefill(const float* q, int M, int n_heads, int n_kv_heads, int head_size, const void* k_cache,
                         const void* v_cache, const int* pos, const int* slot, const int* block_table, int max_pages,
                         int page_size, const int* meta /*device*/, uint8_t* xq_out, uint8_t* xh_out, int out_mcap,
                         float* y_out, int pdl, void* stream);

/* same job on tcgen05 (head_size 128): S = Q K^T and O += P V as UMMA with the score / probability / output tiles in
 * tensor memory, 128 query rows x one head per CTA, 128 tokens per step, lazy softmax rescaling */
int dl_attention_prefill_tc(const float* q, int M, int n_heads, int n_kv_heads, int head_size, const void* k_cache,
                            const void* v_cache, const int* pos, const int* slot, const int* block_table, int max_pages,
                            int page_size, const int* meta /*device*/, uint8_t* xq_out, uint8_t* xh_out, int out_mcap,
                            float* y_out, int n_cache_pages /*pages in k_cache / v_cache: bounds of the TMA tensor maps*/,
                            int pdl, void* stream);

/* ---- sampling helpers --------------------------------------------------------------------- */
/* argmax over logits[m][0..vocab) -> tokens[m]  (reference: Sampler::sample temperature==0,
 * src/tokenizer.cpp:472-502) */
int dl_argmax(const float* logits, int M, int vocab, int ld, int* tokens_out,
              const uint32_t* flags /*NULL: no wait*/, int n_flags, const uint32_t* step_counter,
              uint32_t flag_mul, uint32_t flag_add, int* pos_inc /*NULL or pos[m] += 1*/,
              int* history /*NULL or [cap][M] ring indexed by *step_counter-1*/, int history_cap,
              void* scratch /*zero-initialised, dl_argmax_scratch_bytes(M)*/, int pdl, void* stream);
long long dl_argmax_scratch_bytes(int M);
/* Greedy pick without any device allocation (this entry point has no counterpart in the reference).
 * For callers that are not allowed to call cudaMalloc -- plugins living inside someone else's memory
 * budget, serving loops that ban allocator calls after start-up, hosts that only hand out raw
 * pointers.  Same result as dl_argmax(), but the scratch it needs lives in a fixed 17 KB array inside
 * the library (held by every GPU that loads it, until exit), so the caller passes only logits and the
 * output.  M <= 64 rows, larger M is rejected.  The first call on a device resolves the array's
 * address and should happen before any stream capture; calls on one device share the array, so calls
 * that can overlap on the GPU race -- keep them ordered. */
int dl_argmax_noalloc(const float* logits, int M, int vocab, int ld, int* tokens_out, void* stream);
/* block the stream until flags[i] >= *step_counter*flag_mul+flag_add for i<n (logits gather to the
 * sampling rank; reference: SYNC_NODE_SLICES_EXCEPT_ROOT, src/llm.cpp:432) */
int dl_wait_flags(const uint32_t* flags, int n, const uint32_t* step_counter, uint32_t flag_mul,
                  uint32_t flag_add, void* stream);

/* ---- persistent decode step ----------------------------------------------------------------
 * One cooperative kernel runs a whole forward pass for 1..4 rows (decode): embedding, every layer's
 * norm -> qkv(+RoPE, KV write) -> paged attention -> wo -> norm -> w1|w3 (SiLU*mul) -> w2, final norm,
 * logits, greedy pick.  Phases are separated by device-side grid barriers instead of kernel
 * boundaries and the Q40T weight stream is prefetched across them, so the per-kernel fixed latency
 * of the launch sequence is paid once per token.  Tensor-parallel partial sums travel as Q80 LL
 * words ({4 x int8 | fp16 scale, epoch} in one 8-byte store) written straight into every peer's
 * exchange buffer; the next phase polls the words themselves (no fence, counter or flag).
 * (reference: the whole per-token step list of NnExecutor::forward, src/nn/nn-executor.cpp:134-163,
 * over the graph of src/llm.cpp:126-438; SYNC_NODE_SLICES + OP_MERGE_ADD, src/nn/nn-network.cpp:537-569) */
#define DL_MEGA_MAX_ROWS 4
typedef struct {
    const uint8_t* qkv;               /* Q40T [q_dim + 2 kv_dim][dim]                      */
    const uint8_t* wo;                /* Q40T [dim][q_dim]                                 */
    const uint8_t* w13;               /* Q40T [2 hid][dim], rows interleaved (gate, up)    */
    const uint8_t* w2;                /* Q40T [dim][hid]                                   */
    const float* rms_att;
    const float* rms_ffn;
    void* k_cache;                    /* fp16 [n_pages][n_kv_heads][32][head_size]         */
    void* v_cache;
} DlDecodeLayer;

typedef struct {
    /* model, per-rank slice sizes */
    int n_layers, dim, q_dim, kv_dim, hid, vocab_loc, vocab, head_size, n_heads, n_kv_heads;
    int rank, world;
    float eps;
    int act_gelu;
    int q80_sync;                     /* must be 1: the wo / w2 partial sums travel as Q80 words */
    const DlDecodeLayer* layers;      /* device array [n_layers] */
    const float* rms_final;
    const uint8_t* wcls;              /* Q40T [vocab_loc][dim] */
    const float* embedding;           /* [vocab][dim] f32 */
    const float* rope_freqs;          /* [head_size/2] */
    /* batch */
    int M, row0;                      /* rows; logits / tokens are produced for rows row0..M-1 */
    int* tokens;                      /* [M] */
    int* pos;                         /* [M] */
    const int* slot;                  /* [M] */
    const int* block_table;           /* [n_slots][max_pages] */
    int max_pages;
    /* activations and scratch of this rank (device) */
    float* x[2];                      /* residual stream [4][dim], ping-pong */
    float* q;                         /* [4][q_dim] */
    float* h;                         /* [4][hid] */
    float* att_o;                     /* [4][n_heads][max_splits][head_size] */
    float* att_ml;                    /* [4][n_heads][max_splits][2] */
    int max_splits;                   /* capacity of the split dimension above (<= 32) */
    int att_splits;                   /* split-KV grid: CTAs per (row, kv head) */
    float* cta_best;                  /* [grid][4][2] greedy candidates */
    unsigned* grid_bar;               /* 64 zero-initialised words: [0] arrival counter, [1] exit counter (left at zero by the
                                       * kernel), [16, 32) release flags of the merge-add CTAs (epochs, only grow) */
    uint32_t* step_counter;
    /* tensor-parallel exchange: rank p's buffers as mapped into this process (p < world) */
    uint8_t* ll_att[DL_MAX_PEERS];    /* [world src][4][dim/32][9] 8-byte LL words */
    uint8_t* ll_ffn[DL_MAX_PEERS];
    uint8_t* ll_tok[DL_MAX_PEERS];    /* [world src][4][2] LL words: best logit, its index */
    float* logits[DL_MAX_PEERS];      /* [rows][vocab] on every rank (only logits[rank] when world == 1) */
    uint32_t* flag_logits[DL_MAX_PEERS]; /* word to release on rank p once this rank's logits slice is there */
    /* output */
    int greedy;                       /* 1: pick argmax in the kernel (logits are not written) */
    int* tokens_out;                  /* [M - row0] (greedy) */
    int advance;                      /* 1: tokens[m] = pick, pos[m] += 1 (device-resident loop) */
    int* history;                     /* NULL or [cap][M] ring indexed by step - 1 */
    int history_cap;
    unsigned long long timeout_ns;    /* spin limit before the kernel traps (0 = none) */
    uint32_t* debug;                  /* NULL or a word in mapped pinned host memory: 0x80000000 | wait code << 16 | CTA
                                       * is stored there before a time-out trap (device memory is lost with the context) */
    unsigned long long* trace;        /* NULL or [n_layers * 32 + 32] globaltimer stamps of CTA trace_cta */
    int trace_cta;
    int dbg;                          /* timing probes (results are garbage): 1 = matmul phases skip the arithmetic, 2 = they do not wait for the weights */
    int pf_dist;                      /* pipeline chunks (<= 9 KB) the L2 prefetch of the weight stream runs ahead of the
                                       * shared-memory ring; 0 = no L2 prefetch */
} DlDecodeArgs;

/* 0 if the shapes fit the persistent kernel (shared memory, register-resident row, head size), else -1
 * with the reason in dl_last_error(); callers fall back to the multi-kernel launch sequence. */
int dl_decode_mega_supported(const DlDecodeArgs* args /*host*/, int grid);
int dl_decode_mega(const DlDecodeArgs* args /*host*/, int grid /*CTAs, <= SM count*/, void* stream);
/* work split of one matmul phase for CTA `cta` of `grid`: rows [r0, r1) and the number of pipeline
 * chunks they take (host helper, used by tests) */
int dl_decode_mega_plan(int d, int n, int align, int grid, int cta, int* r0, int* r1, int* n_chunks);

/* ---- peer memory (symmetric buffers over NVLink) ------------------------------------------ */
int dl_ipc_alloc(void** dev_ptr, long long bytes, uint8_t* handle_out /*64 B host*/);
int dl_ipc_open(const uint8_t* handle /*64 B host*/, void** dev_ptr);
int dl_ipc_close(void* dev_ptr);
int dl_ipc_free(void* dev_ptr);
/* all ranks: flag barrier over peer memory (one CTA) */
int dl_peer_barrier(uint32_t* const* flags_on_peers /*host array*/, uint32_t* my_flags, int rank, int world,
                    uint32_t epoch, void* stream);

#ifdef __cplusplus
}
#endif
#endif
