/*
 * kx.h -- C ABI of libkx_b200.so, the B200 (sm_100a) implementation of the hot
 * path of mrankitvish/ai-agent-kubectl.
 *
 * The reference's own boundary for this path is Python-level (a LangChain
 * Runnable in the `llm` slot of `template | llm | KubectlOutputParser()`,
 * app.py:118); it has no FFI.  This header is what a binding of that path would
 * load: plain pointers and sizes, no torch types.  INTEGRATION.md shows the
 * ctypes stub a maintainer of the reference would add.
 *
 * Conventions: every function returns 0 on success and a negative KX_E* code on
 * failure; kx_last_error() returns a thread-local message for the last failure.
 * Pointers named d_* are device pointers on the current CUDA device, h_* are host
 * pointers.  `stream` is a cudaStream_t passed as void* (NULL = default stream).
 * Device entry points only enqueue work; they do not synchronise.
 */
#ifndef KX_H_
#define KX_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KX_OK            0
#define KX_EINVAL       -1   /* bad argument */
#define KX_ECUDA        -2   /* a CUDA runtime call failed */
#define KX_ENOMEM       -3
#define KX_ETOOBIG      -4   /* exceeds a stated limit */

const char *kx_last_error(void);
int kx_version(void);

/* ------------------------------------------------------------------------- *
 * String path: validator + parser + sanitiser + cache-key hash
 * ------------------------------------------------------------------------- */

/* verdict codes: why is_safe_kubectl_command (app.py:72-88) returns False */
#define KX_SAFE          0   /* the reference returns True                        */
#define KX_BAD_PREFIX    1   /* app.py:75-77  does not start with "kubectl "      */
#define KX_METACHAR      2   /* app.py:79-81  ; && || ` $ ( ) < >                 */
#define KX_BAD_QUOTING   3   /* app.py:83-87  shlex.split raised ValueError       */

/* flags */
#define KX_FLAG_PARSE    1u  /* validate as KubectlOutputParser.parse does
                                (app.py:95-104): strip, drop a ```...``` fence,
                                strip again, then validate                        */

/*
 * kx_validate_hash -- batched replacement for, per string s of a CSR batch:
 *     verdict[i]  = reason code of is_safe_kubectl_command(s)       app.py:72-88
 *                   (with KX_FLAG_PARSE: of the command parse() would validate,
 *                    app.py:95-104)
 *     span[2i..]  = byte range [begin,end) inside s of that command (what
 *                   parse() returns when verdict is KX_SAFE)
 *     hash[i]     = hash(sanitize_query(s)) as CPython computes it with
 *                   PYTHONHASHSEED=0 -- the key hash TTLCache's dict uses for
 *                   cache.get / cache[...] at app.py:312,322 (sanitize_query is
 *                   app.py:60-68)
 *     san, san_len= sanitize_query(s) itself, UTF-8, written at d_san + off[i]
 *                   (never longer than s), and its byte length
 * d_data holds the UTF-8 bytes of all strings back to back; string i is
 * d_data[d_off[i] .. d_off[i+1]).  d_off has n+1 entries, d_off[0] need not be 0.
 * Any of the five output pointers may be NULL to skip that output.  Strings must
 * be shorter than 2^31-16 bytes.  Input that is not valid UTF-8 is processed by
 * a fixed, documented rule rather than rejected.
 * Scratch: the first call on a (device, stream) pair with a larger n than before
 * allocates 4 bytes per string of library-owned device memory (cudaMalloc, which
 * synchronises); later calls reuse it.  Calls on one stream are ordered; calls on
 * different streams use different scratch and may overlap.
 * The kernel reads d_data in aligned 16-byte units, so the bytes up to the next
 * 16-byte boundary after the last string must be readable (true of any CUDA
 * allocation).
 */
int kx_validate_hash(const uint8_t *d_data, const int64_t *d_off, int64_t n, uint32_t flags,
                     uint8_t *d_verdict, int64_t *d_hash, int32_t *d_span,
                     uint8_t *d_san, int32_t *d_san_len, void *stream);

/*
 * kx_validate_hash_noalloc -- kx_validate_hash for callers that cannot allocate device
 * memory when they call (no counterpart in the reference; it exists for hosts whose
 * allocator already owns all of HBM, or whose request loop cannot afford the device-wide
 * synchronisation of cudaMalloc).  Same arguments and results as kx_validate_hash, but
 * the scratch queue is a fixed array inside the library, so:
 *   - n may be at most KX_NOALLOC_MAX_STRINGS per call (KX_ETOOBIG otherwise);
 *   - the array (about 4 MB) is resident on every GPU that loads the library;
 *   - all calls share it: calls that overlap on one GPU race, so serialise them;
 *   - no cudaMalloc, cudaFree or pool allocation happens in this call;
 *   - the first call on each device looks up the array's device address; make that
 *     call before any stream capture starts.
 */
#define KX_NOALLOC_MAX_STRINGS (1 << 20)
int kx_validate_hash_noalloc(const uint8_t *d_data, const int64_t *d_off, int64_t n, uint32_t flags,
                             uint8_t *d_verdict, int64_t *d_hash, int32_t *d_span,
                             uint8_t *d_san, int32_t *d_san_len, void *stream);

/*
 * kx_validate_hash_host -- same contract with HOST buffers: copies the batch to
 * `device` in chunks over two streams (copy of chunk k+1 overlaps the kernel on
 * chunk k), runs kx_validate_hash and copies the requested outputs back.
 * Blocks until the outputs are in host memory.  This is the call a Python
 * binding of the reference would make with the strings it already holds
 * (app.py:102 and app.py:305,312).  Staging buffers are cached per device and
 * released by kx_release_host_staging().  h_off[0] must be 0.
 */
int kx_validate_hash_host(const uint8_t *h_data, const int64_t *h_off, int64_t n, uint32_t flags,
                          uint8_t *h_verdict, int64_t *h_hash, int32_t *h_span,
                          uint8_t *h_san, int32_t *h_san_len, int device);
void kx_release_host_staging(void);

/* ------------------------------------------------------------------------- *
 * Generator.  The reference has no code for these: its
 * `llm` slot is ChatOpenAI (app.py:108-117), a remote service; BASELINE.json's
 * north_star defines the kernels.  What they replace is that slot.
 * ------------------------------------------------------------------------- */

/* epilogues of kx_gemm_bf16 */
#define KX_EPI_BF16      0   /* out bf16 = acc (+ bias[feature])                       */
#define KX_EPI_RESID     1   /* out fp32 += acc   (residual stream)                    */
#define KX_EPI_SWIGLU    2   /* features are (gate, up) pairs: out bf16[f/2] = silu(g)*u */
#define KX_EPI_F32       3   /* out fp32 = acc    (logits)                             */
#define KX_EPI_ARGMAX    5   /* lm_head + greedy pick without materialising the logits: out = int2 [parts][ldo],
                                (max logit bits, lowest feature index attaining it) per token and part.
                                swap = 1 (batch <= 256): 4 parts per 128-feature tile; swap = 0: 1 part per
                                256-feature tile.  kx_argmax_parts reduces the parts to one id per token. */

/*
 * kx_gemm_bf16 -- D[a_rows, b_rows] = A[a_rows, k] * B[b_rows, k]^T on tcgen05 tensor
 * cores (bf16 inputs with k contiguous, row pitches lda / ldb elements, fp32
 * accumulation in TMEM), with a fused epilogue.
 *   swap = 0 (prefill): A = activations, B = weights; out[i * ldo + j], feature = j.
 *   swap = 1 (decode) : A = weights, B = activations (batch <= any); out[j * ldo + i],
 *                       feature = i.  Every weight byte is read once.
 * bias (fp32, per feature) may be NULL.  splits > 1 cuts k across CTAs (decode, few
 * feature tiles); partial sums are reduced in library scratch (see kx_release_scratch).
 * k % 8 == 0, ldo % 8 == 0, operands 16-byte aligned.
 */
int kx_gemm_bf16(const void *d_a, int64_t a_rows, int64_t lda, const void *d_b, int64_t b_rows, int64_t ldb,
                 int k, int epilogue, int swap, void *d_out, int64_t ldo, const float *d_bias, int splits,
                 void *stream);

/*
 * kx_rmsnorm -- the decoder's RMSNorm as a stand-alone op (k5; defined by BASELINE.json's north_star, no reference code):
 * d_y bf16 [M, H] = bf16( x * (1 / sqrt(mean(x^2) + eps)) * g ) for the rows of d_x fp32 [*, H] (row i of the output reads
 * row d_row_map[i] of d_x, or row i when d_row_map is NULL); d_g bf16 [H].  splits > 0 (M <= 2048 only): the rows first
 * receive the pending split-K partial sums of the preceding projection, x += d_part[0] + d_part[1] + ... (slice s at
 * d_part + s * split_stride, added in slice order) and are written back -- what the engine does for decode batches
 * of 256 and more.
 */
int kx_rmsnorm(float *d_x, const void *d_g, void *d_y, int32_t M, int32_t H, float eps, const int32_t *d_row_map,
               const float *d_part, int32_t splits, int64_t split_stride, void *stream);

/* Second stage of KX_EPI_ARGMAX: d_parts int2 [n_parts][ld] -> d_ids int32 [B]; ties go to the lowest id. */
int kx_argmax_parts(const void *d_parts, int32_t n_parts, int32_t ld, int32_t B, int32_t *d_ids, void *stream);

/*
 * kx_attn_decode -- one decode step of attention over the paged KV cache, as the
 * engine runs it: for sequence b the new token sits at position pos[b]; its q heads and k head are
 * rotated (rope[pos][hd/2] = (cos, sin), rotate-half pairs), k and v are appended to the cache, and
 * out[b, head, :] = softmax(q K^T / sqrt(hd)) V over keys 0 .. pos[b] (bf16 in/out, fp32 softmax).
 *   d_qkv          [B, qkv_ld] bf16 rows laid out (q heads | k heads | v heads), not yet rotated
 *   d_pages        [num_pages][2 (k, v)][n_kv][16 tokens][hd] bf16
 *   d_block_table  [B (+1), max_pages]: page of tokens 16 i .. 16 i + 15 of each sequence
 *   prefix_len     > 0: the first prefix_len keys (a multiple of 16, >= 64) of EVERY sequence are the
 *                  same pages, listed in block-table row prefix_row; they are read once per 64 query
 *                  rows instead of once per sequence (partials in library scratch, kx_release_scratch)
 *   max_T          upper bound of pos[b] + 1 (sizes the shared-memory rounds)
 * No reference counterpart (the reference calls a remote LLM, app.py:117).
 */
int kx_attn_decode(const void *d_qkv, int32_t qkv_ld, void *d_out, int32_t out_ld, void *d_pages,
                   const int32_t *d_block_table, int32_t max_pages, const int32_t *d_pos, const float *d_rope,
                   int32_t B, int32_t max_T, int32_t n_q, int32_t n_kv, int32_t hd, int32_t prefix_len,
                   int32_t prefix_row, void *stream);

/*
 * Generation engine: a decoder-only transformer (Llama / Qwen2 family) with a paged
 * bf16 KV cache, greedy decoding (temperature 0, app.py:109).  This is what
 * B200ChatModel (the occupant of the reference's `llm` slot, app.py:117-118) drives.
 * All weight pointers are device pointers owned by the caller and must outlive the
 * engine; layouts:
 *   embed   bf16 [vocab, hidden]         lm_head bf16 [vocab, hidden] (= embed when tied)
 *   norm    bf16 [hidden]                rope    fp32 [max_pos, head_dim/2, 2] (cos, sin)
 *   per layer: ln1, ln2 bf16 [hidden]; wqkv bf16 [(n_q + 2 n_kv) head_dim, hidden]
 *   (q heads, then k heads, then v heads); bqkv fp32 [same rows] or NULL;
 *   wo bf16 [hidden, n_q head_dim]; wgu bf16 [2 inter, hidden] with rows interleaved
 *   (gate_0, up_0, gate_1, up_1, ...); wd bf16 [hidden, inter].
 */
typedef struct kx_model_desc {
    int32_t hidden, layers, n_q, n_kv, head_dim, inter, vocab;
    int32_t max_pos;        /* rows of the rope table                                  */
    float   eps;
    int32_t max_batch;      /* sequences resident at once                              */
    int32_t max_seq;        /* prompt + generated tokens per sequence                  */
    int32_t prefill_chunk;  /* max tokens per prefill pass (0 = default 32768)         */
    /* tensor parallelism: with tp_size > 1 the head counts, `inter` and
     * `vocab` above are this rank's SHARD (column-parallel wqkv / wgu, row-parallel wo / wd,
     * vocabulary-parallel lm_head); `hidden` is full.  vocab_offset = first vocabulary id of
     * the shard.  The collectives are supplied with kx_engine_set_collectives.             */
    int32_t tp_size, tp_rank, vocab_offset;
} kx_model_desc;

typedef struct kx_layer_weights {
    const void *ln1, *wqkv;
    const float *bqkv;
    const void *wo, *ln2, *wgu, *wd;
} kx_layer_weights;

typedef struct kx_model_weights {
    const void *embed, *lm_head, *norm;
    const float *rope;
    const kx_layer_weights *layers;   /* host array of `layers` entries */
} kx_model_weights;

typedef struct kx_gen_stats {
    float prefill_ms;      /* enqueue -> first token of every sequence (device time)  */
    float decode_ms;       /* all remaining steps                                      */
    int32_t decode_steps;
    int32_t launches;      /* kernels launched by this call (graph replays counted)    */
    int64_t h2d_bytes;     /* bytes this call copied host -> device (tokens, positions) */
    int64_t d2h_bytes;     /* ... and device -> host (the generated ids)               */
} kx_gen_stats;

typedef struct kx_engine kx_engine;

int kx_engine_create(const kx_model_desc *desc, const kx_model_weights *weights, kx_engine **out);

/*
 * Tensor-parallel engines reduce the row-parallel partial sums (two all-reduces of
 * [tokens, hidden] fp32 per layer) and merge the per-shard argmax (one all-gather of
 * [batch] (value, id) pairs per step) through caller-supplied NCCL entry points:
 *   comm       an ncclComm_t spanning the tp_size ranks
 *   allreduce  address of ncclAllReduce, allgather address of ncclAllGather
 * They are enqueued on the engine's stream (and captured into its decode graph).
 */
int kx_engine_set_collectives(kx_engine *e, void *comm, void *allreduce, void *allgather);
void kx_engine_destroy(kx_engine *e);
int64_t kx_engine_device_bytes(const kx_engine *e);

/*
 * Fused tensor-parallel all-reduce for decode (tp_size 2): the row-parallel GEMMs (wo, wd) push their
 * partial sums into the peer GPU's memory from their epilogue and add the peer's, instead of calling
 * NCCL.  Each rank exports two CUDA IPC handles (2 x 64 bytes: receive buffer, flags), the host swaps
 * them between the ranks, and each rank imports its peer's.  Without this set-up the engine uses the
 * NCCL path of kx_engine_set_collectives for every token count.
 */
int kx_engine_p2p_export(kx_engine *e, void *handles);
int kx_engine_p2p_import(kx_engine *e, const void *peer_handles);
int64_t kx_engine_p2p_launches(const kx_engine *e);   /* GEMMs enqueued with the fused all-reduce so far */

/* Prefill attention for head_dim 64 (with or without a cached prefix) runs on tcgen05 tensor cores (csrc/kx_attn_tc.cu);
 * other shapes, and every shape with KX_ATTN_TC=0 in the environment at engine creation, take the mma.sync
 * kernel.  Returns how many attention launches took the tcgen05 kernel so far. */
int64_t kx_engine_attn_tc_launches(const kx_engine *e);
/* Diagnostic: with KX_TRACE=1 in the environment at creation every instrumented kernel of a decode step records
 * globaltimer values into its slot: [0] earliest CTA start, [1] earliest return from griddepcontrol.wait,
 * [2] latest end of the main loop, [3] latest CTA end, [4] latest end of the split-K exchange (ns), [5..7] spare.
 * Copies the slots of the LAST step (launch order) to h_slots (8 x uint64 each) and returns how many; kx_engine_trace_name gives the launch's label. */
int kx_engine_trace(kx_engine *e, uint64_t *h_slots, int32_t max_slots);
const char *kx_engine_trace_name(const kx_engine *e, int32_t i);

/*
 * kx_engine_set_prefix -- prefill `n` tokens (a multiple of 16; host int32) ONCE and keep their
 * KV pages; every later kx_engine_generate treats its prompts as continuing that prefix
 * (positions start at n, the pages are shared by all sequences through the block table).
 * This is the prefix cache for the constant part of the reference's prompt template
 * (app.py:50-57) and for BASELINE config 5's system prompt.  n = 0 removes the prefix.
 */
int kx_engine_set_prefix(kx_engine *e, const int32_t *h_tokens, int32_t n, void *stream);

/*
 * kx_engine_set_eos -- end-of-sequence token id (negative: none, the default).  Decode steps of at most 256
 * sequences on a single-GPU replica stop being launched once every sequence has emitted it (checked every 8 steps);
 * the remaining positions of h_out are filled with the id.  The reference's LLM ends a completion the same way
 * (the OpenAI service stops at its end-of-text token; app.py:117 only sees the finished text).
 */
int kx_engine_set_eos(kx_engine *e, int32_t eos_id);

/*
 * kx_engine_generate -- prefill `batch` prompts of `prompt_len` tokens each (host int32
 * [batch, prompt_len]) and decode `n_new` tokens greedily; h_out receives int32
 * [batch, n_new].  d_logits (device fp32 [n_new, batch, vocab]) may be NULL; when given,
 * the logits that produced every token are kept (for the parity tests) and the decode
 * loop is launched kernel by kernel instead of as a CUDA graph.  Blocks until done.
 */
int kx_engine_generate(kx_engine *e, const int32_t *h_tokens, int32_t batch, int32_t prompt_len, int32_t n_new,
                       int32_t *h_out, float *d_logits, kx_gen_stats *stats, void *stream);

/*
 * Continuous batching (kx_engine_session_*).  Replaces, in the reference, the way /kubectl-command serves requests:
 * each one is awaited on its own as it arrives (app.py:183-186) and is cancelled on its own when its deadline passes
 * (app.py:189-191).  The engine's max_batch sequence SLOTS are admitted, stepped and released independently, so a
 * request that arrives while others are decoding joins at the next step boundary instead of waiting for their batch
 * to end, and a cancelled or finished request frees its slot (and KV pages) at once.  Single-GPU replicas only.
 * One host thread drives a session; kx_engine_generate and kx_engine_set_prefix refuse to run while slots are live
 * and end the session (open it again afterwards).
 *
 *   open     every slot free.
 *   admit    prefill n prompts of prompt_len tokens each (host int32 [n, prompt_len]) into slots slot0 .. slot0+n-1
 *            (free ones: the caller keeps the slot map) and choose each one's first token; the slots then produce up
 *            to max_new tokens, fewer if kx_engine_set_eos's id appears.  Blocks until the prefill is done; the
 *            decoding slots keep their state across it.  stats (may be NULL): prefill_ms, launches, h2d_bytes.
 *   step     n_steps decode steps over slots 0 .. n_slots-1 (n_slots = highest live slot + 1), enqueued on the engine's
 *            stream as CUDA graph replays; returns without waiting.  Free and finished slots idle in place.
 *   poll     waits for the enqueued steps; h_count[s] = tokens slot s has produced (-1: free), h_finished[s] = 1 once
 *            it has ended (both host int32 [max_batch]).
 *   read     the first n tokens of the outputs of slots slot0 .. slot0+n_slots-1 (host int32 [n_slots, n]).
 *   release  slots slot0 .. slot0+n-1 become free (finished, or cancelled mid-sequence).
 *   live     slots admitted and not yet released.
 */
int kx_engine_session_open(kx_engine *e);
int kx_engine_session_admit(kx_engine *e, int32_t slot0, int32_t n, const int32_t *h_tokens, int32_t prompt_len, int32_t max_new,
                            kx_gen_stats *stats);
int kx_engine_session_step(kx_engine *e, int32_t n_steps, int32_t n_slots);
int kx_engine_session_poll(kx_engine *e, int32_t *h_count, int32_t *h_finished);
int kx_engine_session_read(kx_engine *e, int32_t slot0, int32_t n_slots, int32_t *h_out, int32_t n);
int kx_engine_session_release(kx_engine *e, int32_t slot0, int32_t n);
int32_t kx_engine_session_live(const kx_engine *e);

/* frees the device scratch the library caches per (device, stream): the queue
 * kx_validate_hash uses to hand strings from its fast pass to its general pass
 * (4 bytes per string of the largest batch seen on that stream) */
void kx_release_scratch(void);

/* number of kernels this library has launched in this process (for bench.py) */
int64_t kx_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* KX_H_ */
