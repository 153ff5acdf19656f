/* m2f_b200.h -- C ABI of the B200-native M2FNet hot path (libm2f_b200.so).
 *
 * The reference (iosonopersia/Multimodal-Emotion-Recognition) has no FFI of its own: the
 * boundary the hot path sits behind is the Python class `M2FNet` in src/model.py.  Every
 * entry point below cites the reference arithmetic it replaces (paths relative to the
 * reference repo; "torch/" = the PyTorch the reference executes through).  INTEGRATION.md
 * shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *  - plain C types only; every pointer is a DEVICE pointer unless named host_*;
 *  - every function returns 0 (M2F_OK) or a negative error code; the message is available
 *    from m2f_last_error() (thread-local).  No C++ exception crosses the boundary;
 *  - all work is enqueued on `stream` (a cudaStream_t passed as void*), never synchronises,
 *    never allocates or frees device memory: the caller owns every buffer;
 *  - unsupported shapes are errors, never silent fallbacks.  There is no CPU path.
 *
 * "planes": a matrix stored as P in {1,2,3} bf16 planes whose sum is the value
 * (P = 1: bf16 throughput mode; P = 3: ~24 mantissa bits, the fp32-parity mode in which each
 * tensor-core GEMM runs the 6 significant plane products with fp32 accumulation).
 */
#ifndef M2F_B200_H
#define M2F_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M2F_ABI_VERSION 2

#define M2F_OK 0
#define M2F_ERR_ARG (-1)         /* bad argument / shape */
#define M2F_ERR_UNSUPPORTED (-2) /* valid request outside the kernels' limits */
#define M2F_ERR_CUDA (-3)        /* CUDA runtime / driver error */
#define M2F_ERR_ALIGN (-4)       /* pointer or stride alignment */

/* dropout call sites (src/model.py:57,61,73,85,97,111-131).  mod: 0 audio, 1 text */
#define M2F_SITE_ENC(mod, t, l, k) (0x1000u * ((mod) + 1u) + 0x100u * (t) + 0x10u * (l) + (k))
#define M2F_SITE_MOD(mod, k) (0x1000u * ((mod) + 1u) + 0xF00u + (k))
#define M2F_SITE_FAM(i, k) (0x3000u + 0x10u * (i) + (k))
#define M2F_SITE_CLS 0x4000u

typedef struct {
    void* ptr;            /* bf16, plane 0 */
    int64_t rows, cols;   /* logical extent */
    int64_t ld;           /* row stride in elements; multiple of 8 */
    int64_t plane_stride; /* elements between consecutive planes */
    int32_t planes;       /* 1..3 */
    int32_t _pad;
} m2f_planes_t;

typedef struct {
    int32_t a_k0, b_k0, klen; /* contraction segment: A[:, a_k0:a_k0+klen] . B[:, b_k0:b_k0+klen]^T */
} m2f_kseg_t;

/* Epilogue applied to each accumulator element v = sum_k A[m,k] B[n,k], in this order:
 *   v += bias[n]; if relu: v = max(v,0); if gate: v = gate[m,n] > 0 ? v*gate_scale : 0;
 *   if drop_p > 0: v = keep(seed,step,site, m*drop_ld+n) ? v/(1-p) : 0; v += resid[m,n];
 *   if accumulate: v += out_f32[m,n];  then stored to out_f32 and/or out_planes; colsum[n] += v (atomic). */
typedef struct {
    const float* bias;
    int32_t relu;
    int32_t accumulate;
    const void* gate; /* bf16 plane 0 of a saved activation */
    int64_t gate_ld;
    float gate_scale;
    float drop_p;
    uint64_t seed, step;
    uint32_t site;
    int32_t _pad;
    int64_t drop_ld;
    const float* resid;
    int64_t resid_ld;
    float* out_f32;
    int64_t out_f32_ld;
    m2f_planes_t out_planes; /* ptr == NULL: none */
    float* colsum;           /* nullable: colsum[n] += sum over rows m < M of the stored value (bias gradients) */
} m2f_epilogue_t;

int m2f_abi_version(void);
const char* m2f_last_error(void);

/* Number of kernels this library has launched in this process so far (bench.py's `gpu_launches`). */
int64_t m2f_launch_count(void);

/* Dropout randomness from device memory: while `seed_step` (device pointer to {uint64 seed, uint64 step}) is non-NULL,
 * every dropout-carrying launch made from the calling host thread reads its (seed, step) from there when the kernel
 * runs, and ignores the by-value pair -- what lets a captured CUDA graph be replayed with a new step counter.
 * NULL restores the by-value behaviour. */
int m2f_set_rng_source(const uint64_t* seed_step);

/* Device capability check: returns M2F_OK only on an sm_100 device. */
int m2f_check_device(void);

/* ---- K1: tcgen05 GEMM.  D[M,N] = sum over segments and plane products of A . B^T.
 * Replaces every nn.Linear / in-projection of the model (src/model.py:9,68,80,90-98;
 * torch/nn/functional.py:5833-5873; torch/nn/modules/transformer.py:980-982) and, with
 * a_mn_major / b_mn_major, their autograd backward (dgrad: B = W stored [N_contract, K_out];
 * wgrad: A = dY stored [M_contract, N_out], B = X stored [M_contract, K_out]).
 * x_mn_major = 0: operand stored [rows = M or N, cols = K]; 1: stored [rows = K, cols = M or N]. */
int m2f_gemm(const m2f_planes_t* A, int a_mn_major, const m2f_planes_t* B, int b_mn_major,
             const m2f_kseg_t* segs, int nsegs, int64_t M, int64_t N,
             const m2f_epilogue_t* epi, void* stream);

/* fp32 [rows, cols] (row stride src_ld) -> bf16 planes (weights shadow copy, model inputs). */
int m2f_split_planes(const float* src, int64_t src_ld, const m2f_planes_t* dst, void* stream);
/* dst = relu(src) on planes (first fusion layer's concat operand, src/model.py:16-17) */
int m2f_relu_planes(const m2f_planes_t* src, const m2f_planes_t* dst, void* stream);
/* out[n] += sum_m src[m,n]  (bias gradients) */
int m2f_colsum_planes(const m2f_planes_t* src, float* out, void* stream);

/* ---- K3: LayerNorm (+ residual / outer skip).  nn.LayerNorm eps 1e-5, biased variance
 * (torch/nn/modules/transformer.py:977-982 norm1/norm2; :550-551 final norm; src/model.py:107,119).
 *   y = LN(x) * gamma + beta (+ skip);  stats[2*m] = mean, stats[2*m+1] = rstd.
 *   y_f32 (nullable) and y_planes (nullable ptr) receive y; the planes copy optionally gets
 *   dropout (site, p) applied -- the model-level dropout before the projection (src/model.py:111,123). */
int m2f_ln_fwd(const float* x, int64_t x_ld, const float* gamma, const float* beta,
               const float* skip, int64_t skip_ld, int64_t M, int64_t D,
               float* y_f32, int64_t y_ld, const m2f_planes_t* y_planes,
               float drop_p, uint64_t seed, uint64_t step, uint32_t site,
               float* stats, void* stream);

/* LayerNorm backward.  g = dL/dy [M,D] fp32;  x, stats from the forward.
 *   dx = rstd * (gamma*g - mean(gamma*g) - xhat * mean(gamma*g*xhat));
 *   dgamma += sum_m g*xhat; dbeta += sum_m g;
 *   dx_f32 (nullable) receives dx; dx_planes (nullable ptr) receives dropmask(dx) (the mask of the
 *   dropout that sat between the producing Linear and the residual add, regenerated from
 *   (seed, step, site)); dbias (nullable) += column sums of dropmask(dx). */
int m2f_ln_bwd(const float* g, int64_t g_ld, const float* x, int64_t x_ld, const float* gamma,
               const float* stats, int64_t M, int64_t D,
               float* dx_f32, int64_t dx_ld, const m2f_planes_t* dx_planes,
               float drop_p, uint64_t seed, uint64_t step, uint32_t site,
               float* dgamma, float* dbeta, float* dbias, void* stream);

/* ---- K2: masked multi-head attention, one CTA per (dialogue, head).
 * P = softmax(q k^T / sqrt(hd) with -inf on padded KEYS), O = dropout(P) v
 * (torch/nn/functional.py:6607-6690; fusion layers: q,v from text, k from audio, src/model.py:14).
 * q,k,v,o: planes views of [B*S, H*hd] (cols = H*hd; they may alias one packed buffer through
 * different base pointers and a common ld).  key_pad: uint8 [B,S], 1 = padded.
 * Dropout element of probability (b, h, query i, key j): ((b*H + h)*S + i) * S8 + j of `site`, S8 = S rounded up to 8.
 * Limits (M2F_ERR_UNSUPPORTED beyond).  Head sizes 16 / 32 / 64 / 96 / 128 run on tensor cores: S <= 64 with the whole problem on
 * chip; 64 < S <= 4096 (head sizes 32 and up) tiled, forward and backward, with or without dropout; inference with one plane, head
 * sizes 64 / 96 / 128 and S <= 256 on tcgen05.  Any other head size runs a generic kernel: S <= 256 and a shared-memory footprint
 * (forward 2 S hd, backward 4 S hd + 2 S^2 floats) within 227 KB. */
int m2f_attn_fwd(const m2f_planes_t* q, const m2f_planes_t* k, const m2f_planes_t* v,
                 const uint8_t* key_pad, int64_t B, int64_t S, int32_t H,
                 float drop_p, uint64_t seed, uint64_t step, uint32_t site,
                 const m2f_planes_t* o, void* stream);
/* backward: recomputes P; d_o planes in, dq/dk/dv planes out.  dbq/dbk/dbv (nullable, fp32 [H*hd]) += column sums of
 * dq/dk/dv over all B*S rows: the in_proj_bias gradient slices (torch/nn/functional.py:5833-5873). */
int m2f_attn_bwd(const m2f_planes_t* q, const m2f_planes_t* k, const m2f_planes_t* v,
                 const m2f_planes_t* d_o, const uint8_t* key_pad, int64_t B, int64_t S, int32_t H,
                 float drop_p, uint64_t seed, uint64_t step, uint32_t site,
                 const m2f_planes_t* dq, const m2f_planes_t* dk, const m2f_planes_t* dv,
                 float* dbq, float* dbk, float* dbv, void* stream);

/* ---- K4: cross entropy, ignore_index = -1, label smoothing eps, optional class weights
 * (src/train.py:48-50,229).  logits fp32 [M,C] (row stride ld), target int64 [M].
 * loss_out[0] = loss, loss_out[1] = sum of weights over valid rows (scratch: 4 floats, zeroed inside).
 * dlogits (nullable) fp32 [M, dl_ld] receives grad_scale * dL/dlogits; dl_planes likewise. */
int m2f_ce_fwd_bwd(const float* logits, int64_t ld, const int64_t* target, const float* class_w,
                   int64_t M, int32_t C, float eps, float grad_scale, float* loss_out,
                   float* dlogits, int64_t dl_ld, const m2f_planes_t* dl_planes, void* stream);

/* ---- K4s: the same cross entropy for callers that cannot allocate.  No counterpart in the reference: added
 * (code under CUDA graph capture, serving loops and harnesses that forbid cudaMalloc).
 * Arguments as m2f_ce_fwd_bwd, except that loss_out is 2 floats of pure OUTPUT (loss_out[0] = loss, loss_out[1] = sum
 * of weights over valid rows) and nothing the caller passes is used as scratch: the running sums and one log-sum-exp
 * per row live in a file-scope device array of the library (about 1 MB, held by every GPU that loads the library until
 * the process exits).  Limit: M <= M2F_CE_STATIC_MAX_ROWS, beyond it M2F_ERR_UNSUPPORTED.  The call performs no
 * cudaMalloc / cudaFree.  All calls on one device share the array (it keeps the last call's contents; what must start
 * at zero is cleared on `stream` at the start of the call): calls that overlap on the GPU race -- serialise them. */
#define M2F_CE_STATIC_MAX_ROWS 262144
int m2f_ce_fwd_bwd_static(const float* logits, int64_t ld, const int64_t* target, const float* class_w,
                          int64_t M, int32_t C, float eps, float grad_scale, float* loss_out,
                          float* dlogits, int64_t dl_ld, const m2f_planes_t* dl_planes, void* stream);


/* ---- model-level API: the whole M2FNet forward / backward as one call each ------------------------------
 * m2f_model_cfg_t mirrors the `model:` block of src/config.yaml:31-56 (what M2FNet.__init__ reads,
 * src/model.py:24-56) plus the build-side knobs (planes, ffn_dim = nn.TransformerEncoderLayer's default 2048). */
typedef struct {
    int32_t enabled, dim, n_head, n_transformers, n_layers;
} m2f_modality_cfg_t;

typedef struct {
    m2f_modality_cfg_t audio, text;
    int32_t fam_enabled, fam_dim, fam_heads, fam_layers;
    int32_t cls_hidden, cls_out, cls_layers;
    float dropout;
    int32_t ffn_dim;
    int32_t planes; /* planes of everything the FORWARD multiplies (weights shadow, saved activations): 1 = plain bf16,
                     * 2 = 16-bit mantissa, 3 = fp32-parity mode */
    int32_t planes_bwd; /* planes of the BACKWARD's GEMM / attention operands: the gradients, and the leading planes_bwd
                         * planes of the saved activations and weights.  0 = same as `planes`; must be <= planes.
                         * (planes, planes_bwd) = (2, 1) is the "bf16x2f" training mode: ReLU gates and softmax
                         * weights are decided by a 16-bit-mantissa forward -- what the gradient error at bf16 is
                         * dominated by (profiles/r2_bf16_protocol_cpu.json) -- while both backward products per
                         * Linear stay single bf16 products: 5 plane products per Linear instead of 9 (bf16x2). */
} m2f_model_cfg_t;

typedef struct m2f_model m2f_model_t;
typedef struct m2f_plan m2f_plan_t;

/* Parameter layout (host only; works without a GPU).  All parameters live in ONE fp32 arena: 2-D weights
 * first, then every 1-D tensor from m2f_model_arena_1d_offset() on.  Names are the reference state_dict
 * names; the shared final norm appears once, as `{m}_encoders.0.norm.*`.
 * Invalid enable-flag combinations return M2F_ERR_ARG with the reference's message (src/model.py:32-35). */
int m2f_model_create(const m2f_model_cfg_t* cfg, m2f_model_t** out);
int m2f_model_destroy(m2f_model_t* m);
int64_t m2f_model_num_params(const m2f_model_t* m);
/* cols == 0: 1-D tensor of length rows.  offset in floats from the arena base. */
int m2f_model_param_info(const m2f_model_t* m, int64_t i, char* name_buf, int32_t name_buf_len, int64_t* offset,
                         int64_t* rows, int64_t* cols);
int64_t m2f_model_arena_floats(const m2f_model_t* m);
int64_t m2f_model_arena_1d_offset(const m2f_model_t* m);
/* bf16-plane shadow copy of the 2-D weights (GEMM B operands, TMA-readable) + fused bias copies */
int64_t m2f_model_shadow_bytes(const m2f_model_t* m);
int m2f_model_refresh_shadow(const m2f_model_t* m, const float* params, void* shadow, void* stream);

/* A plan fixes the batch geometry (B dialogues x S padded utterances) and owns the workspace layout and the
 * cached TMA descriptors.  training_plan = 1 keeps every activation the backward needs. */
int m2f_plan_create(m2f_model_t* m, int64_t B, int64_t S, int32_t training_plan, m2f_plan_t** out);
int m2f_plan_destroy(m2f_plan_t* p);
int64_t m2f_plan_workspace_bytes(const m2f_plan_t* p);
/* M2FNet.forward (src/model.py:102-145).  text [B,S,Dt] / audio [B,S,Da] fp32 contiguous, key_pad uint8 [B,S]
 * (1 = padding), logits fp32 [B,S,C] contiguous.  dropout_on = model.training (p from the config). */
int m2f_plan_forward(m2f_plan_t* p, const float* params, const void* shadow, void* workspace, const float* text,
                     const float* audio, const uint8_t* key_pad, int32_t dropout_on, uint64_t seed, uint64_t step,
                     float* logits, void* stream);
/* Fused criterion on the logits of the last forward (src/train.py:48-50,229): writes loss_out[0..3] and keeps
 * grad_scale * dL/dlogits in the workspace for m2f_plan_backward(dlogits = NULL). */
int m2f_plan_loss(m2f_plan_t* p, void* workspace, const float* logits, const int64_t* target, const float* class_w,
                  float eps, float grad_scale, float* loss_out, void* stream);
/* Backward of the last forward.  dlogits fp32 [B,S,C] (from autograd) or NULL (use m2f_plan_loss's).
 * Every parameter gradient is WRITTEN (not accumulated) into the fp32 arena `grads` (same layout as params). */
int m2f_plan_backward(m2f_plan_t* p, const float* params, const void* shadow, void* workspace, const float* dlogits,
                      float* grads, void* stream);

/* Data-parallel variant: `cb(user, lo, hi, ready)` is called on the host, in enqueue order, each time at least
 * `bucket_floats` gradients have accumulated: the arena range [lo, hi) (floats) is final once the work enqueued so far
 * on the CUDA stream `ready` has completed (the plan runs the backward on several internal streams; `ready` is the one
 * ordered after the producers of this range, not necessarily `stream`).  The 2-D weights complete from the top of the
 * arena downwards (classifier, fusion layers, text stack, audio stack), the 1-D region
 * [m2f_model_arena_1d_offset, arena_floats) last.  The caller launches its NCCL all-reduce of that range behind
 * `ready` from the callback, so the exchange overlaps the rest of the backward.  When the
 * call returns, everything (all internal streams) has been joined back into `stream`. */
typedef void (*m2f_bucket_cb)(void* user, int64_t lo, int64_t hi, void* ready);
int m2f_plan_backward_bucketed(m2f_plan_t* p, const float* params, const void* shadow, void* workspace,
                               const float* dlogits, float* grads, int64_t bucket_floats, m2f_bucket_cb cb, void* user,
                               void* stream);

/* ---- gradient exchange over NVLink / NVSwitch peer memory -------------------------------------
 * The reference trains on one device (src/train.py:20); the data-parallel path adds exactly one exchange between
 * loss.backward() (src/train.py:230) and optimizer.step() (src/train.py:231): the all-reduce of the gradient arena.
 * Here it is one kernel over a *symmetric* gradient buffer: the same allocation mapped on every GPU of the node
 * (peer_base[r] = rank r's copy as seen from this GPU, own copy included) and, where the fabric offers it, through an
 * NVSwitch multicast object (mc_base, else NULL).  The host obtains both mappings from its symmetric-memory allocator
 * (torch.distributed._symmetric_memory: buffer_ptrs / multicast_ptr); the library only consumes the addresses.
 * `signal_off_bytes` locates M2F_XCHG_MAX_CTAS x M2F_XCHG_MAX_RANKS uint32 barrier words inside the buffer, behind
 * the gradients, zeroed once before the first call (and after any aborted one). */
#define M2F_XCHG_MAX_RANKS 16
#define M2F_XCHG_MAX_CTAS 64
#define M2F_XCHG_NO_EXIT_BARRIER 1
#define M2F_XCHG_WIRE_BF16 2 /* the range travels as bf16 (needs wire_off_bytes, the multicast mapping, 8-float alignment) */
#define M2F_XCHG_SIGNAL_BYTES (M2F_XCHG_MAX_CTAS * M2F_XCHG_MAX_RANKS * 4)
typedef struct {
    void* mc_base;                        /* multicast mapping of the buffer (multimem.ld_reduce / multimem.st), or NULL */
    void* peer_base[M2F_XCHG_MAX_RANKS];  /* unicast mapping of every rank's buffer; entries >= world are ignored */
    int64_t signal_off_bytes;
    int32_t rank, world;
    int64_t wire_off_bytes;               /* bf16 staging area inside the buffer (2 bytes per gradient float, element i at
                                           * wire_off_bytes + 2 i), behind the signal words; 0 = none (fp32 wire only) */
    int32_t wire_bf16;                    /* m2f_plan_set_exchange: exchange the ranges of the backward in bf16 */
    int32_t _pad;
} m2f_peer_group_t;

/* In-place all-reduce of the float range [lo, hi) of the symmetric buffer: on return of the *kernel* every rank holds
 * scale * sum over ranks.  Two-shot: rank r reduces slice r (in the switch with multicast, by world peer loads without)
 * and writes it to every rank.  Stream-ordered; every rank must issue the same sequence of calls (same lo, hi, ctas)
 * and have one GPU to itself -- the kernels of the ranks wait for each other.  ctas <= 0: one CTA per 128 KB of a rank's slice, at most 16 (enough to saturate NVLink).
 * flags: M2F_XCHG_NO_EXIT_BARRIER skips the closing handshake -- allowed on every call of a sequence on one stream except the
 * last one, whose handshake then covers the whole sequence (consumers must wait for that last kernel).
 * With world == 1 it degenerates to an in-place scale (used by the single-GPU tests of the slicing arithmetic).
 * M2F_XCHG_WIRE_BF16: the range is packed to bf16 into the staging area, all-reduced there (multimem.ld_reduce
 * .acc::f32.v4.bf16x2: fp32 accumulation in the switch, half the NVLink bytes) and unpacked back to fp32 -- three launches, the
 * middle one with both handshakes; the result carries two bf16 roundings (relative 2^-9).  Ranges that are not whole
 * 8-float vectors, groups without multicast or staging area, and world == 1 silently use the fp32 wire. */
int m2f_grad_allreduce_peer(const m2f_peer_group_t* g, int64_t lo, int64_t hi, float scale, int32_t ctas, int32_t flags,
                            void* stream);

/* Tuning knob: CTA shape of the exchange kernel -- threads x unroll x 16 B of loads in flight per CTA; one of
 * (512,8) [default], (512,4), (256,8), (256,16), (1024,4), (128,16).  Process-wide; every rank must set the same. */
int m2f_exchange_tune(int32_t threads, int32_t unroll);

/* The exchange as part of the backward: after this call m2f_plan_backward_bucketed(..., bucket_floats > 0, cb = NULL, ...)
 * all-reduces every finished gradient range itself -- m2f_grad_allreduce_peer on the plan's ready stream, captured into
 * the backward's CUDA graph, so a data-parallel step needs no host work per range -- and the caller's stream continues
 * only after the last range has been exchanged.  `grads` of that call must be this rank's copy of the symmetric buffer
 * (group->peer_base[group->rank]).  The group is copied.  group == NULL detaches.  m2f_plan_exchange_ranges: number of
 * ranges the last such backward exchanged (identical on every rank). */
int m2f_plan_set_exchange(m2f_plan_t* p, const m2f_peer_group_t* group, float scale, int32_t ctas);
int64_t m2f_plan_exchange_ranges(const m2f_plan_t* p);

/* Debug aid: the exchange kernel keeps a ring of the last M2F_XCHG_TRACE_RECORDS launches on this device, 6 x uint64 each:
 * %globaltimer (ns) at entry, after the first barrier, after the data phase, at exit; lo; floats.  Copies the ring to
 * out_records (host, M2F_XCHG_TRACE_RECORDS * 6 values; record i of the process sits at slot i % RECORDS) and returns
 * the number of launches so far.  Synchronises the device. */
#define M2F_XCHG_TRACE_RECORDS 512
int64_t m2f_exchange_trace(uint64_t* out_records);

/* Adam with L2 weight decay folded into the gradient (torch.optim.Adam semantics, src/train.py:56,231), one
 * launch over the arena; `t` is the 1-based step count.  grad_scale multiplies g first (DDP averaging). */
int m2f_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                  float beta1, float beta2, float eps, float weight_decay, int64_t t, float grad_scale, void* stream);

/* The same update for a model's arena, fused with the refresh of the bf16 weight shadow (m2f_model_refresh_shadow):
 * each updated 2-D weight is split into its planes while still in registers, saving one pass over the parameters.
 * Bit-identical to m2f_adam_step over the whole arena followed by m2f_model_refresh_shadow. */
int m2f_model_adam_step(const m2f_model_t* m, float* params, const float* grads, float* exp_avg, float* exp_avg_sq,
                        void* shadow, float lr, float beta1, float beta2, float eps, float weight_decay, int64_t t,
                        float grad_scale, void* stream);

/* The optimizer behind the exchange.  With an exchange attached (m2f_plan_set_exchange), what follows the backward's last
 * kernel is the backlog of the exchange stream and then optimizer.step() (src/train.py:231) over the whole arena.  After
 * m2f_plan_set_optimizer the backward itself enqueues the Adam update (and the refresh of the bf16 weight shadow) of every
 * exchanged gradient range behind its compute, each gated only by the exchange call after its own -- so the update of the early
 * ranges overlaps the exchange of the late ones -- all captured in the backward's CUDA graph; step() then only counts.  Same
 * arithmetic, bit for bit, as m2f_model_adam_step after the backward.  The hyper-parameters live in device memory
 * (M2F_ADAM_HYPER_FLOATS floats, 16-byte aligned) so that a replayed graph sees the current step count / learning rate: call
 * m2f_adam_hyper on the same stream before each such backward (t = 1-based count of the update about to be applied).
 * Without an exchange attached the setting has no effect.  exp_avg == NULL detaches. */
#define M2F_ADAM_HYPER_FLOATS 8
int m2f_plan_set_optimizer(m2f_plan_t* p, float* exp_avg, float* exp_avg_sq, const float* hyper_dev);
int m2f_adam_hyper(float* hyper_dev, float lr, float beta1, float beta2, float eps, float weight_decay, int64_t t,
                   float grad_scale, void* stream);

/* ---- either side of the step ---------------------------------------------------
 * Batch assembly on the device, replacing Dataset.__getitem__ + collate_fn + apply_padding
 * (src/dataset.py:32-65,71-89; src/utils.py:15-31).  The embedding tables stay resident in HBM: text_table
 * fp32 [n_rows, Dt] (row stride text_ld), audio_table fp32 [n_rows, Da], labels int64 [n_rows] (one per CSV row).
 * Dialogue d owns utt_rows[dlg_offsets[d] .. dlg_offsets[d+1]) = its table rows sorted by Utterance_ID.
 * batch_dialogues int64 [B] picks the dialogues.  Outputs (contiguous): text_out [B,S,Dt] and audio_out [B,S,Da]
 * padded with 0, emotion_out int64 [B,S] padded with -1, mask_out uint8 [B,S] = (emotion == -1).  All pointers
 * are device pointers; the caller guarantees S >= every selected dialogue's length (longer ones are truncated). */
int m2f_collate_dialogues(const float* text_table, int64_t text_ld, int32_t Dt, const float* audio_table,
                          int64_t audio_ld, int32_t Da, const int64_t* labels, const int64_t* utt_rows,
                          const int64_t* dlg_offsets, const int64_t* batch_dialogues, int64_t B, int64_t S,
                          float* text_out, float* audio_out, int64_t* emotion_out, uint8_t* mask_out, void* stream);

/* Per-batch accuracy_score and f1_score(average='weighted') of src/test.py:65-71 / src/train.py:262-267 without the
 * per-batch device->host copy: argmax over the C logits of each row (first maximum), confusion int32 [C*C]
 * (overwritten; row = true class, column = predicted) over rows whose target != ignore_index, then
 * sums[0] += accuracy, sums[1] += weighted F1, sums[2] += 1 (double[3], accumulated across calls: the reference
 * averages the per-batch scores over the batches).  pred (nullable) int64 [M] receives the argmax.  C <= 32. */
int m2f_batch_metrics(const float* logits, int64_t ld, const int64_t* target, int64_t M, int32_t C,
                      int64_t ignore_index, int32_t* confusion, double* sums, int64_t* pred, void* stream);

/* test helper: mask[i] = 1 if element i of `site` is kept, for i in [0, n) */
int m2f_dropout_mask(float p, uint64_t seed, uint64_t step, uint32_t site, int64_t n, uint8_t* mask,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* M2F_B200_H */
