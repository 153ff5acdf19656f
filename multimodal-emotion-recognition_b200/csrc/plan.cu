// Model-level executor: the whole M2FNet forward (src/model.py:102-145) and its backward as one C call each.
//
// Host-side only (no kernels here except the shadow refresh): it owns the parameter-arena layout, the
// bf16-plane shadow copy of the weights, the per-(B,S) workspace layout and a cache of TMA descriptors,
// and enqueues the kernels of gemm_sm100.cu / attention.cu / layernorm.cu / loss.cu on the caller's stream.
//
// Data flow (M = B*S rows, kept [B,S,D] row-major end to end -- the reference's permutes,
// head split/merge copies and torch.cat calls never materialise):
//   * the residual stream, LayerNorm inputs and every gradient w.r.t. the residual stream stay fp32;
//   * everything a GEMM reads (activations and the weight shadow) is bf16 planes (P = 1 or 3);
//   * concat operands are column halves of one buffer (fusion `u_i`, classifier input `zc`).
#include <string>
#include <unordered_map>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "gemm.cuh"

namespace m2f {

// ------------------------------------------------------------------------------------------ layout types
enum Base { WS = 0, SHADOW = 1 };

struct Buf {  // a matrix inside the workspace or the shadow, by byte offset
    int64_t off = -1;
    int64_t rows = 0, cols = 0, ld = 0, stride = 0;
    int planes = 0;  // 0: fp32
    Base base = WS;
};

struct ParamInfo {
    std::string name;
    int64_t off, rows, cols;
};

struct Lin {       // one nn.Linear (or a fused group of in-projection rows)
    Buf w;         // shadow planes [out, in]
    int64_t w_off; // arena offset of the fp32 weight (gradient destination)
    int64_t b_off; // arena offset of the fp32 bias
    int out, in;
};

struct EncLayer {
    Lin qkv, wo, w1, w2;
    int64_t g1, be1, g2, be2;  // LayerNorm weight / bias arena offsets
    // saved activations
    Buf qkv_pl, o_pl, y1, st1, x1, x1_pl, h_pl, y2, st2, x2, x2_pl;
};

struct Modality {
    int enabled = 0, idx = 0, D = 0, H = 0, T = 0, L = 0;
    std::vector<EncLayer> layers;  // T * L
    int64_t gf = 0, bef = 0;       // shared final norm
    Lin proj;
    Buf in_pl;                     // planes of the raw input
    std::vector<Buf> xout, xout_pl, stf;  // per transformer: x_t (fp32), its planes, final-norm stats
};

struct FusionLayer {
    Lin qv, wo, lin;  // qv: rows [Wq; Wv] of in_proj_weight fused in the shadow
    int64_t inw_off, inb_off;  // arena offsets of in_proj_weight / in_proj_bias
    Buf bqv;                   // fp32 shadow copy [bq; bv]
    Buf qv_pl, xo_pl, u_pl;    // saved: [q | v], attention output, relu([x | t])
};

struct SEntry {  // shadow refresh work item
    int64_t src_off;
    int rows, cols, src_ld;
    int64_t dst_off;
    int dst_ld;
    int64_t stride;
    int planes;  // 0: fp32 copy
};

}  // namespace m2f

using namespace m2f;

struct m2f_model {
    m2f_model_cfg_t cfg;
    std::vector<ParamInfo> params;
    int64_t arena_floats = 0, off_1d = 0;
    int64_t shadow_bytes = 0;
    std::vector<SEntry> sentries;
    Modality mod[2];
    std::vector<FusionLayer> fam;
    Lin wk_all;   // fused K projections of all fusion layers [F*Df, Df]
    Buf bk_all;   // fp32 shadow copy of their biases
    std::vector<Lin> cls;
    int Df = 0, cls_in = 0;
    bool both = false;

    int64_t add_param(const std::string& name, int64_t rows, int64_t cols, int64_t& cursor) {
        cursor = round_up(cursor, 8);  // 32 bytes: float4 access, and ranges of whole 16-byte bf16 vectors for the exchange's bf16 wire
        params.push_back({name, cursor, rows, cols});
        const int64_t off = cursor;
        cursor += rows * (cols ? cols : 1);
        return off;
    }
    Buf add_shadow_planes(int64_t src_off, int rows, int cols, int src_ld) {
        Buf b;
        b.base = SHADOW;
        b.rows = rows;
        b.cols = cols;
        b.ld = round_up(cols, 8);
        b.stride = round_up((int64_t)rows * b.ld, 8);
        b.planes = cfg.planes;
        shadow_bytes = round_up(shadow_bytes, 256);
        b.off = shadow_bytes;
        shadow_bytes += b.stride * 2 * cfg.planes;
        sentries.push_back({src_off, rows, cols, src_ld, b.off, (int)b.ld, b.stride, cfg.planes});
        return b;
    }
    // rows appended to an existing shadow matrix (fused projections)
    void add_shadow_rows(const Buf& dst, int row0, int64_t src_off, int rows, int cols, int src_ld) {
        sentries.push_back({src_off, rows, cols, src_ld, dst.off + (int64_t)row0 * dst.ld * 2, (int)dst.ld, dst.stride, cfg.planes});
    }
    Buf alloc_shadow_planes(int rows, int cols) {
        Buf b;
        b.base = SHADOW;
        b.rows = rows;
        b.cols = cols;
        b.ld = round_up(cols, 8);
        b.stride = round_up((int64_t)rows * b.ld, 8);
        b.planes = cfg.planes;
        shadow_bytes = round_up(shadow_bytes, 256);
        b.off = shadow_bytes;
        shadow_bytes += b.stride * 2 * cfg.planes;
        return b;
    }
    Buf alloc_shadow_f32(int n) {
        Buf b;
        b.base = SHADOW;
        b.rows = 1;
        b.cols = n;
        b.ld = n;
        b.planes = 0;
        shadow_bytes = round_up(shadow_bytes, 256);
        b.off = shadow_bytes;
        shadow_bytes += (int64_t)n * 4;
        return b;
    }
    void add_shadow_f32(const Buf& dst, int col0, int64_t src_off, int n) {
        sentries.push_back({src_off, 1, n, n, dst.off + (int64_t)col0 * 4, n, 0, 0});
    }
};

struct TmapKey {
    const void* ptr;
    int64_t rows, cols, ld, stride;
    int planes, box_rows;
    bool operator==(const TmapKey& o) const {
        return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && stride == o.stride && planes == o.planes &&
               box_rows == o.box_rows;
    }
};
struct TmapKeyHash {
    size_t operator()(const TmapKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        h = h * 1000003u ^ (size_t)k.rows;
        h = h * 1000003u ^ (size_t)k.cols;
        h = h * 1000003u ^ (size_t)k.ld;
        h = h * 1000003u ^ (size_t)(k.planes * 1024 + k.box_rows);
        return h;
    }
};

struct EpiKey {
    const void *out_f32, *out_pl, *gate, *resid;
    int64_t out_f32_ld, out_pl_ld, out_pl_stride, gate_ld, resid_ld;
    int M, N, out_pl_n;
    bool operator==(const EpiKey& o) const { return std::memcmp(this, &o, sizeof(EpiKey)) == 0; }
};
struct EpiKeyHash {
    size_t operator()(const EpiKey& k) const {
        size_t h = 1469598103934665603ull;
        const unsigned char* b = reinterpret_cast<const unsigned char*>(&k);
        for (size_t i = 0; i < sizeof(EpiKey); ++i) h = (h ^ b[i]) * 1099511628211ull;
        return h;
    }
};
struct EpiEntry {
    EpiMaps maps;
    int epi_tma;
};

struct m2f_plan {
    m2f_model* m;
    int64_t B, S, M;
    int training;
    int64_t ws_bytes = 0;
    // per-plan copies of the model's layer tables (their activation Bufs are plan-specific)
    Modality mod[2];
    std::vector<FusionLayer> fam;
    // workspace buffers that are not per-layer
    Buf key_pad;                            // uint8 [M] copy of the padding mask (re-read by the backward)
    Buf zc_pl, t0_pl, kall_pl;
    std::vector<Buf> cls_h;                 // classifier hidden activations
    Buf dl_pl;                              // dlogits planes [M, 8-padded]
    Buf logits_copy;                        // unused placeholder
    // backward scratch
    // per modality (the two stacks run concurrently): fp32 chain buffers and two alternating sets (layer parity) of the
    // operand buffers that the weight-gradient stream reads while the data-gradient chain has moved on
    struct StackScratch {
        Buf gG, gA, gB;                     // fp32 [M, D]
        Buf dx2[2], dx1[2], dh[2], do_[2], dqkv[2];
    } ss[2];
    Buf dcls_a, dcls_b;                     // classifier hidden gradients (ping-pong)
    Buf dz[3], dxo[2], dattn[2], dqv[2];    // fusion backward (rotating: read by the weight-gradient stream)
    Buf dkall_pl, dproj_pl, dprojt_pl, gt_f32, ga_f32;
    // concurrency: aux[0] = audio stack chain, aux[1] = weight gradients of classifier / fusion / text stack,
    // aux[2] = weight gradients of the audio stack; a ring of events orders work across them
    cudaStream_t aux[3] = {nullptr, nullptr, nullptr};
    cudaStream_t ready = nullptr;            // gradient buckets are reported as "final on this stream"
    cudaEvent_t wg_ev[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // weight gradients of (modality, layer parity) done
    cudaEvent_t wgf_ev[3] = {nullptr, nullptr, nullptr};                // ... of fusion layer f % 3
    std::vector<cudaEvent_t> evs;
    size_t ev_next = 0;
    int use_streams = 1;
    // CUDA graphs: the body of the forward / backward is captured once (on `cap`, with its forks onto aux[]) and replayed
    // into the caller's stream; whatever depends on per-call pointers (inputs, logits, dlogits) stays outside.
    struct GraphKey {
        const void *params = nullptr, *shadow = nullptr, *ws = nullptr, *grads = nullptr, *opt_m = nullptr;
        float p_used = -1.f;
        int flags = 0;
        int64_t bucket_floats = 0;
        bool operator==(const GraphKey& o) const {
            return params == o.params && shadow == o.shadow && ws == o.ws && grads == o.grads && opt_m == o.opt_m && p_used == o.p_used &&
                   flags == o.flags && bucket_floats == o.bucket_floats;
        }
    };
    struct GraphSlot {
        cudaGraphExec_t exec = nullptr;
        GraphKey key;
        long long kernels = 0;
        std::vector<std::pair<int64_t, int64_t>> marks;  // gradient ranges reported through external events (data parallel)
        int64_t xchg_ranges = 0;                         // exchange kernels captured in this graph
        void reset() {
            if (exec) cudaGraphExecDestroy(exec);
            exec = nullptr;
        }
    } gf, gb;
    std::vector<cudaEvent_t> bucket_evs;     // recorded by event nodes of the backward graph, one per reported range
    std::vector<cudaEvent_t> stack_evs;      // "weight gradients of this stack down to here are enqueued" (interleaved stack reports)
    // gradient exchange inside the backward (m2f_plan_set_exchange): each finished range is all-reduced by a kernel on `ready`
    bool xchg_on = false;
    m2f_peer_group_t xchg_group = {};
    float xchg_scale = 1.f;
    int xchg_ctas = 0;
    int64_t xchg_ranges = 0;                 // ranges exchanged by the last backward
    // optimizer behind the exchange (m2f_plan_set_optimizer): once the backward's compute has finished, each exchanged gradient
    // range is updated as soon as IT is final, while the exchange of the later ranges is still running
    bool opt_on = false;
    float *opt_m = nullptr, *opt_v = nullptr;
    const float* opt_hp = nullptr;           // device: AdamArgs, rewritten by m2f_adam_hyper before each backward
    std::vector<cudaEvent_t> xchg_evs;       // recorded on `ready` behind each exchange call
    cudaStream_t cap = nullptr;
    unsigned long long* rng_dev = nullptr;   // {seed, step} read by the dropout kernels of the captured graphs
    int use_graph = 1;
    Buf text_in, audio_in;                   // fp32 staging of the inputs (graph mode: the captured kernels read these)

    int ensure_streams() {
        if (aux[0]) return M2F_OK;
        {   // Priorities (M2F_STREAM_PRIO=0: all equal): the data-gradient chains are the critical path, the weight gradients
            // fill the SMs they leave free; the exchange launched behind `ready` is a few CTAs that should start as soon as
            // their range is final.  Kernel nodes captured from these streams keep the priority.
            int lo_prio = 0, hi_prio = 0;
            M2F_CUDA(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
            const char* e = std::getenv("M2F_STREAM_PRIO");
            const bool prio = !(e && e[0] == '0');
            const int chain_prio = prio ? (hi_prio + lo_prio) / 2 : lo_prio;  // numerically lower = more urgent
            M2F_CUDA(cudaStreamCreateWithPriority(&aux[0], cudaStreamNonBlocking, chain_prio));
            M2F_CUDA(cudaStreamCreateWithPriority(&aux[1], cudaStreamNonBlocking, lo_prio));
            M2F_CUDA(cudaStreamCreateWithPriority(&aux[2], cudaStreamNonBlocking, lo_prio));
            M2F_CUDA(cudaStreamCreateWithPriority(&ready, cudaStreamNonBlocking, hi_prio));
            M2F_CUDA(cudaStreamCreateWithPriority(&cap, cudaStreamNonBlocking, chain_prio));
        }
        M2F_CUDA(cudaMalloc(&rng_dev, 2 * sizeof(unsigned long long)));
        evs.resize(128);
        for (auto& e : evs) M2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j) M2F_CUDA(cudaEventCreateWithFlags(&wg_ev[i][j], cudaEventDisableTiming));
        for (int j = 0; j < 3; ++j) M2F_CUDA(cudaEventCreateWithFlags(&wgf_ev[j], cudaEventDisableTiming));
        return M2F_OK;
    }
    void release_streams() {
        for (auto& e : evs) cudaEventDestroy(e);
        evs.clear();
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j)
                if (wg_ev[i][j]) { cudaEventDestroy(wg_ev[i][j]); wg_ev[i][j] = nullptr; }
        for (int j = 0; j < 3; ++j)
            if (wgf_ev[j]) { cudaEventDestroy(wgf_ev[j]); wgf_ev[j] = nullptr; }
        gf.reset();
        gb.reset();
        for (auto& e : bucket_evs) cudaEventDestroy(e);
        bucket_evs.clear();
        for (auto& e : stack_evs) cudaEventDestroy(e);
        stack_evs.clear();
        for (auto& e : xchg_evs) cudaEventDestroy(e);
        xchg_evs.clear();
        if (rng_dev) { cudaFree(rng_dev); rng_dev = nullptr; }
        if (tl_dev) { cudaFree(tl_dev); tl_dev = nullptr; }
        if (cap) { cudaStreamDestroy(cap); cap = nullptr; }
        if (ready) { cudaStreamDestroy(ready); ready = nullptr; }
        for (int i = 0; i < 3; ++i)
            if (aux[i]) { cudaStreamDestroy(aux[i]); aux[i] = nullptr; }
    }
    // debug timeline (M2F_TIMELINE=<file>): a 1-thread kernel after every op stamps %globaltimer on the op's stream
    bool tl_on = false;
    unsigned long long* tl_dev = nullptr;
    std::vector<std::pair<std::string, int>> tl_labels;
    int64_t tl_calls = 0;
    // state of the last forward
    float p_used = 0.f;
    uint64_t seed = 0, step = 0;
    bool have_loss_grad = false;
    // descriptor cache
    const void* bound_ws = nullptr;
    const void* bound_shadow = nullptr;
    std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> tmaps;
    std::unordered_map<EpiKey, EpiEntry, EpiKeyHash> epimaps;

    Buf alloc_pl(int64_t rows, int64_t cols, int planes) {
        Buf b;
        b.rows = rows;
        b.cols = cols;
        b.ld = round_up(cols, 8);
        b.stride = round_up(rows * b.ld, 8);
        b.planes = planes;
        ws_bytes = round_up(ws_bytes, 256);
        b.off = ws_bytes;
        ws_bytes += b.stride * 2 * planes;
        return b;
    }
    Buf alloc_f32(int64_t rows, int64_t cols) {
        Buf b;
        b.rows = rows;
        b.cols = cols;
        b.ld = round_up(cols, 4);
        b.planes = 0;
        ws_bytes = round_up(ws_bytes, 256);
        b.off = ws_bytes;
        ws_bytes += rows * b.ld * 4;
        return b;
    }
};

namespace m2f {

// ------------------------------------------------------------------------------------------ model layout
static std::string enc_prefix(const char* m, int t, int l) {
    return std::string(m) + "_encoders." + std::to_string(t) + ".layers." + std::to_string(l) + ".";
}

static int build_model(m2f_model* md) {
    const m2f_model_cfg_t& c = md->cfg;
    // the reference's constructor checks (src/model.py:32-35), same messages
    M2F_REQUIRE(c.audio.enabled || c.text.enabled, M2F_ERR_ARG, "At least one of audio and text must be enabled!");
    M2F_REQUIRE(!c.fam_enabled || (c.audio.enabled && c.text.enabled), M2F_ERR_ARG,
                "Fusion Attention Module can only be used with both audio and text enabled!");
    M2F_REQUIRE(c.planes >= 1 && c.planes <= 3, M2F_ERR_ARG, "planes must be 1, 2 or 3");
    M2F_REQUIRE(c.planes_bwd >= 0 && c.planes_bwd <= c.planes, M2F_ERR_ARG, "planes_bwd must be 0 (= planes) or 1..planes");
    M2F_REQUIRE(c.dropout >= 0.f && c.dropout < 1.f, M2F_ERR_ARG, "dropout must be in [0, 1)");
    M2F_REQUIRE(c.ffn_dim > 0 && c.ffn_dim % 8 == 0, M2F_ERR_ARG, "ffn_dim must be a positive multiple of 8");
    M2F_REQUIRE(c.fam_dim > 0 && c.fam_dim % 8 == 0, M2F_ERR_UNSUPPORTED, "FAM.embedding_size must be a positive multiple of 8");
    M2F_REQUIRE(c.cls_hidden > 0 && c.cls_hidden % 8 == 0 && c.cls_out >= 1 && c.cls_out <= 32, M2F_ERR_UNSUPPORTED,
                "CLASSIFIER: hidden_size must be a multiple of 8 and output_size in 1..32");
    const m2f_modality_cfg_t* mc[2] = {&c.audio, &c.text};
    const char* mname[2] = {"audio", "text"};
    for (int i = 0; i < 2; ++i) {
        if (!mc[i]->enabled) continue;
        M2F_REQUIRE(mc[i]->dim > 0 && mc[i]->n_head > 0 && mc[i]->dim % mc[i]->n_head == 0, M2F_ERR_ARG,
                    "embed_dim must be divisible by num_heads");  // torch's assertion text
        M2F_REQUIRE(mc[i]->dim % 4 == 0 && mc[i]->dim <= 1024, M2F_ERR_UNSUPPORTED, "%s embedding_size must be a multiple of 4 and <= 1024",
                    mname[i]);
        M2F_REQUIRE(mc[i]->n_transformers >= 1 && mc[i]->n_transformers <= 15 && mc[i]->n_layers >= 1 && mc[i]->n_layers <= 16,
                    M2F_ERR_UNSUPPORTED, "%s: n_transformers must be 1..15 and n_encoder_layers 1..16", mname[i]);
    }
    if (c.fam_enabled) {
        M2F_REQUIRE(c.fam_heads > 0 && c.fam_dim % c.fam_heads == 0, M2F_ERR_ARG, "embed_dim must be divisible by num_heads");
        M2F_REQUIRE(c.fam_layers >= 1 && c.fam_layers <= 255, M2F_ERR_UNSUPPORTED, "FAM.n_layers must be 1..255");
    }
    md->Df = c.fam_dim;
    md->both = c.audio.enabled && c.text.enabled;
    md->cls_in = md->both ? 2 * md->Df : md->Df;
    const int FF = c.ffn_dim, Df = md->Df;

    // ---- pass 1: 2-D weights
    int64_t cur = 0;
    for (int i = 0; i < 2; ++i) {
        Modality& mo = md->mod[i];
        mo.enabled = mc[i]->enabled;
        mo.idx = i;
        if (!mo.enabled) continue;
        mo.D = mc[i]->dim; mo.H = mc[i]->n_head; mo.T = mc[i]->n_transformers; mo.L = mc[i]->n_layers;
        const int D = mo.D;
        mo.layers.resize((size_t)mo.T * mo.L);
        for (int t = 0; t < mo.T; ++t)
            for (int l = 0; l < mo.L; ++l) {
                EncLayer& e = mo.layers[(size_t)t * mo.L + l];
                const std::string P = enc_prefix(mname[i], t, l);
                e.qkv.w_off = md->add_param(P + "self_attn.in_proj_weight", 3 * D, D, cur); e.qkv.out = 3 * D; e.qkv.in = D;
                e.wo.w_off = md->add_param(P + "self_attn.out_proj.weight", D, D, cur);     e.wo.out = D; e.wo.in = D;
                e.w1.w_off = md->add_param(P + "linear1.weight", FF, D, cur);               e.w1.out = FF; e.w1.in = D;
                e.w2.w_off = md->add_param(P + "linear2.weight", D, FF, cur);               e.w2.out = D; e.w2.in = FF;
            }
        mo.proj.w_off = md->add_param(std::string(mname[i]) + "_proj.weight", Df, D, cur); mo.proj.out = Df; mo.proj.in = D;
    }
    if (c.fam_enabled) {
        md->fam.resize(c.fam_layers);
        for (int f = 0; f < c.fam_layers; ++f) {
            FusionLayer& fl = md->fam[f];
            const std::string P = "fusion_layers." + std::to_string(f) + ".";
            fl.inw_off = md->add_param(P + "multihead_attention.in_proj_weight", 3 * Df, Df, cur);
            fl.wo.w_off = md->add_param(P + "multihead_attention.out_proj.weight", Df, Df, cur); fl.wo.out = Df; fl.wo.in = Df;
            fl.lin.w_off = md->add_param(P + "linear.weight", Df, 2 * Df, cur);                  fl.lin.out = Df; fl.lin.in = 2 * Df;
            fl.qv.out = 2 * Df; fl.qv.in = Df; fl.qv.w_off = fl.inw_off;
        }
    }
    {   // classifier: Linear, (ReLU, Linear) x (n-2), ReLU, Dropout, Linear  (src/model.py:89-100)
        const int n_lin = c.cls_layers > 2 ? c.cls_layers : 2;
        md->cls.resize(n_lin);
        int k = 0, in = md->cls_in;
        for (int j = 0; j < n_lin; ++j) {
            const int idx = (j == 0) ? 0 : (j < n_lin - 1 ? k + 1 : k + 2);
            const int out = (j == n_lin - 1) ? c.cls_out : c.cls_hidden;
            md->cls[j].w_off = md->add_param("output_layer." + std::to_string(idx) + ".weight", out, in, cur);
            md->cls[j].out = out; md->cls[j].in = in;
            md->cls[j].b_off = idx;  // temporarily the Sequential index
            k = (j == 0) ? 1 : k + 2;
            in = out;
        }
    }
    // ---- pass 2: 1-D tensors
    cur = round_up(cur, 64);
    md->off_1d = cur;
    for (int i = 0; i < 2; ++i) {
        Modality& mo = md->mod[i];
        if (!mo.enabled) continue;
        const int D = mo.D;
        for (int t = 0; t < mo.T; ++t)
            for (int l = 0; l < mo.L; ++l) {
                EncLayer& e = mo.layers[(size_t)t * mo.L + l];
                const std::string P = enc_prefix(mname[i], t, l);
                e.qkv.b_off = md->add_param(P + "self_attn.in_proj_bias", 3 * D, 0, cur);
                e.wo.b_off = md->add_param(P + "self_attn.out_proj.bias", D, 0, cur);
                e.w1.b_off = md->add_param(P + "linear1.bias", FF, 0, cur);
                e.w2.b_off = md->add_param(P + "linear2.bias", D, 0, cur);
                e.g1 = md->add_param(P + "norm1.weight", D, 0, cur);
                e.be1 = md->add_param(P + "norm1.bias", D, 0, cur);
                e.g2 = md->add_param(P + "norm2.weight", D, 0, cur);
                e.be2 = md->add_param(P + "norm2.bias", D, 0, cur);
            }
        mo.gf = md->add_param(std::string(mname[i]) + "_encoders.0.norm.weight", D, 0, cur);
        mo.bef = md->add_param(std::string(mname[i]) + "_encoders.0.norm.bias", D, 0, cur);
        mo.proj.b_off = md->add_param(std::string(mname[i]) + "_proj.bias", Df, 0, cur);
    }
    for (int f = 0; f < (int)md->fam.size(); ++f) {
        FusionLayer& fl = md->fam[f];
        const std::string P = "fusion_layers." + std::to_string(f) + ".";
        fl.inb_off = md->add_param(P + "multihead_attention.in_proj_bias", 3 * Df, 0, cur);
        fl.wo.b_off = md->add_param(P + "multihead_attention.out_proj.bias", Df, 0, cur);
        fl.lin.b_off = md->add_param(P + "linear.bias", Df, 0, cur);
    }
    for (auto& L : md->cls) {
        const int idx = (int)L.b_off;
        L.b_off = md->add_param("output_layer." + std::to_string(idx) + ".bias", L.out, 0, cur);
    }
    md->arena_floats = round_up(cur, 64);

    // ---- shadow layout
    for (int i = 0; i < 2; ++i) {
        Modality& mo = md->mod[i];
        if (!mo.enabled) continue;
        for (auto& e : mo.layers) {
            for (Lin* L : {&e.qkv, &e.wo, &e.w1, &e.w2}) L->w = md->add_shadow_planes(L->w_off, L->out, L->in, L->in);
        }
        mo.proj.w = md->add_shadow_planes(mo.proj.w_off, mo.proj.out, mo.proj.in, mo.proj.in);
    }
    if (!md->fam.empty()) {
        const int F = (int)md->fam.size();
        md->wk_all.out = F * Df; md->wk_all.in = Df; md->wk_all.w_off = -1; md->wk_all.b_off = -1;
        md->wk_all.w = md->alloc_shadow_planes(F * Df, Df);
        md->bk_all = md->alloc_shadow_f32(F * Df);
        for (int f = 0; f < F; ++f) {
            FusionLayer& fl = md->fam[f];
            fl.qv.w = md->alloc_shadow_planes(2 * Df, Df);
            md->add_shadow_rows(fl.qv.w, 0, fl.inw_off, Df, Df, Df);                            // Wq
            md->add_shadow_rows(fl.qv.w, Df, fl.inw_off + (int64_t)2 * Df * Df, Df, Df, Df);    // Wv
            md->add_shadow_rows(md->wk_all.w, f * Df, fl.inw_off + (int64_t)Df * Df, Df, Df, Df);  // Wk
            fl.bqv = md->alloc_shadow_f32(2 * Df);
            md->add_shadow_f32(fl.bqv, 0, fl.inb_off, Df);
            md->add_shadow_f32(fl.bqv, Df, fl.inb_off + 2 * Df, Df);
            md->add_shadow_f32(md->bk_all, f * Df, fl.inb_off + Df, Df);
            fl.wo.w = md->add_shadow_planes(fl.wo.w_off, Df, Df, Df);
            fl.lin.w = md->add_shadow_planes(fl.lin.w_off, Df, 2 * Df, 2 * Df);
        }
    }
    for (auto& L : md->cls) L.w = md->add_shadow_planes(L.w_off, L.out, L.in, L.in);
    md->shadow_bytes = round_up(md->shadow_bytes, 256);
    return M2F_OK;
}

// ------------------------------------------------------------------------------------------ shadow refresh
constexpr int SHADOW_CHUNK = 48;
struct ShadowBatch {
    SEntry e[SHADOW_CHUNK];
    int n;
};

__global__ void __launch_bounds__(256)
shadow_refresh_kernel(const float* __restrict__ params, uint8_t* __restrict__ shadow, const ShadowBatch batch) {
    const SEntry& e = batch.e[blockIdx.y];
    const float* src = params + e.src_off;
    if (e.planes == 0) {
        float* dst = reinterpret_cast<float*>(shadow + e.dst_off);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < e.cols; i += gridDim.x * blockDim.x) dst[i] = src[i];
        return;
    }
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(shadow + e.dst_off);
    if ((e.cols & 3) == 0 && (e.src_ld & 3) == 0) {
        const int c4 = e.cols >> 2;
        const int64_t total = (int64_t)e.rows * c4;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = i / c4;
            const int cc = (int)(i - r * c4) * 4;
            const float4 q = *reinterpret_cast<const float4*>(src + r * e.src_ld + cc);
            const float v[4] = {q.x, q.y, q.z, q.w};
            store_planes4(dst, e.stride, e.planes, r * e.dst_ld + cc, v);
        }
    } else {
        const int64_t total = (int64_t)e.rows * e.cols;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = i / e.cols;
            const int cc = (int)(i - r * e.cols);
            store_planes(dst, e.stride, e.planes, r * e.dst_ld + cc, src[r * e.src_ld + cc]);
        }
    }
}

// Adam update of the 2-D weights fused with the refresh of their bf16 shadow: same arithmetic as adam_kernel (optim.cu),
// walked entry by entry so that each updated value is split into planes while it is still in registers.
struct AdamArgs {
    float b1, b2, eps, wd, step_size, inv_sqrt_bc2, grad_scale;
};
__device__ __forceinline__ float adam_one(float& p, float g, float& m, float& v, const AdamArgs& a) {
    const float gr = g * a.grad_scale + a.wd * p;
    m = a.b1 * m + (1.f - a.b1) * gr;
    v = a.b2 * v + (1.f - a.b2) * gr * gr;
    const float denom = sqrtf(v) * a.inv_sqrt_bc2 + a.eps;
    p -= a.step_size * (m / denom);
    return p;
}
// `hp` (nullable): the hyper-parameters in device memory instead of the by-value copy -- what lets the update be captured into
// the backward's CUDA graph and replayed with a new step count / learning rate (m2f_adam_hyper writes them before each launch).
__global__ void __launch_bounds__(256)
adam_shadow_kernel(float* __restrict__ params, const float* __restrict__ grads, float* __restrict__ em, float* __restrict__ ev,
                   uint8_t* __restrict__ shadow, const ShadowBatch batch, const AdamArgs a_val, const AdamArgs* __restrict__ hp) {
    const AdamArgs a = hp ? *hp : a_val;
    const uint64_t pol = l2_evict_first_policy();  // everything here is touched once per step: do not displace the GEMMs' operands
    const SEntry& e = batch.e[blockIdx.y];
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(shadow + e.dst_off);
    if ((e.cols & 3) == 0 && (e.src_ld & 3) == 0 && (e.src_off & 3) == 0) {
        const int c4 = e.cols >> 2;
        const int64_t total = (int64_t)e.rows * c4;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = i / c4;
            const int cc = (int)(i - r * c4) * 4;
            const int64_t idx = e.src_off + r * e.src_ld + cc;
            float4 pp = ldg_stream4(params + idx, pol);
            const float4 gg = ldg_stream4(grads + idx, pol);
            float4 mm = ldg_stream4(em + idx, pol), vv = ldg_stream4(ev + idx, pol);
            float o[4];
            o[0] = adam_one(pp.x, gg.x, mm.x, vv.x, a);
            o[1] = adam_one(pp.y, gg.y, mm.y, vv.y, a);
            o[2] = adam_one(pp.z, gg.z, mm.z, vv.z, a);
            o[3] = adam_one(pp.w, gg.w, mm.w, vv.w, a);
            stg_stream4(params + idx, pp, pol);
            stg_stream4(em + idx, mm, pol);
            stg_stream4(ev + idx, vv, pol);
            store_planes4(dst, e.stride, e.planes, r * e.dst_ld + cc, o);
        }
    } else {
        const int64_t total = (int64_t)e.rows * e.cols;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = i / e.cols;
            const int cc = (int)(i - r * e.cols);
            const int64_t idx = e.src_off + r * e.src_ld + cc;
            float pp = params[idx], mm = em[idx], vv = ev[idx];
            adam_one(pp, grads[idx], mm, vv, a);
            params[idx] = pp; em[idx] = mm; ev[idx] = vv;
            store_planes(dst, e.stride, e.planes, r * e.dst_ld + cc, pp);
        }
    }
}

// the 1-D region (biases, LayerNorm vectors) with device-resident hyper-parameters: same arithmetic as adam_kernel (optim.cu)
__global__ void __launch_bounds__(256)
adam_range_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v, int64_t n4,
                  const AdamArgs* __restrict__ hp) {
    const AdamArgs a = *hp;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 pp = p[i], gg = g[i], mm = m[i], vv = v[i];
        adam_one(pp.x, gg.x, mm.x, vv.x, a);
        adam_one(pp.y, gg.y, mm.y, vv.y, a);
        adam_one(pp.z, gg.z, mm.z, vv.z, a);
        adam_one(pp.w, gg.w, mm.w, vv.w, a);
        p[i] = pp; m[i] = mm; v[i] = vv;
    }
}
__global__ void adam_hyper_kernel(AdamArgs* dst, const AdamArgs a) { *dst = a; }

// ------------------------------------------------------------------------------------------ execution context
__global__ void timeline_stamp_kernel(unsigned long long* slot) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    *slot = t;
}
constexpr int TL_CAP = 4096;

struct Ctx {
    m2f_plan* p;
    const float* params;
    float* grads;
    uint8_t* shadow;
    uint8_t* ws;
    cudaStream_t st;   // the caller's stream: everything forks from and joins back into it
    int P;             // planes of the forward's operands (weights shadow, saved activations)
    cudaStream_t cur = nullptr;  // stream the next kernels are enqueued on

    bool narrow = false;         // prefer half-width GEMM tiles when a launch would fill less than half of the SMs (M2F_NARROW=1)
    cudaStream_t stream() const { return cur ? cur : st; }
    // work enqueued on `to` after this call waits for everything enqueued on `from` so far
    bool forked[3] = {false, false, false};  // side streams that carry work of this call (only those are joined back)
    int dep(cudaStream_t from, cudaStream_t to) {
        if (from == to) return M2F_OK;
        for (int i = 0; i < 3; ++i) {
            if (to == p->aux[i]) forked[i] = true;
            if (from == p->aux[i] && !forked[i]) return M2F_OK;  // nothing was enqueued there
        }
        cudaEvent_t e = p->evs[p->ev_next++ % p->evs.size()];
        M2F_CUDA(cudaEventRecord(e, from));
        M2F_CUDA(cudaStreamWaitEvent(to, e, 0));
        return M2F_OK;
    }
    cudaStream_t side(int i) const { return p->use_streams ? p->aux[i] : st; }
    void stamp(const char* label) {
        if (!p->tl_on || (int)p->tl_labels.size() >= TL_CAP) return;
        cudaStream_t s = stream();
        const int sid = (s == st) ? 0 : (s == p->aux[0] ? 1 : (s == p->aux[1] ? 2 : 3));  // main, audio chain, wgrad, wgrad
        timeline_stamp_kernel<<<1, 1, 0, s>>>(p->tl_dev + p->tl_labels.size());
        p->tl_labels.emplace_back(std::string(label).substr(0, 70), sid);
    }
    // gradient-bucket notification (data-parallel overlap): the 2-D gradients complete top-down in the arena
    int64_t bucket_floats = 0;
    m2f_bucket_cb cb = nullptr;
    void* cb_user = nullptr;
    int64_t hi = 0;

    // Every 2-D gradient at arena offsets >= w has been enqueued.  When a bucket is due, the plan's
    // ready stream is ordered after the weight-gradient streams and handed to the callback: the caller's collective waits on it alone.
    int done_down_to(int64_t w, bool flush = false) {
        if (!cb && !marks && !xchg) return M2F_OK;
        if (collecting >= 0) {  // inside a modality stack: remember the range behind an event on the stack's weight-gradient stream
            if (hi - w >= bucket_floats || (flush && hi > w)) {
                if (stack_ev_used == p->stack_evs.size()) {
                    cudaEvent_t e;
                    M2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    p->stack_evs.push_back(e);
                }
                cudaEvent_t e = p->stack_evs[stack_ev_used++];
                M2F_CUDA(cudaEventRecord(e, collect_stream));
                stack_ranges[collecting].push_back({w, hi, e});
                hi = w;
            }
            return M2F_OK;
        }
        if (hi - w >= bucket_floats || (flush && hi > w)) {
            cudaStream_t ready = p->use_streams ? p->ready : st;
            if (marks || capturing) M2F_TRY(dep(st, ready));  // capture: pulls `ready` into the graph even when no side stream has run yet
            M2F_TRY(dep(side(1), ready));  // the two weight-gradient streams
            M2F_TRY(dep(side(2), ready));
            M2F_TRY(report(w, hi, ready));
            hi = w;
        }
        return M2F_OK;
    }
    // The two modality stacks run side by side and finish together, but are enqueued one after the other: reporting their
    // ranges in enqueue order would make every audio range wait for the whole text stack and leave a stack's worth of
    // exchange behind the end of the backward.  Their ranges are therefore collected per stack, each behind an event on
    // that stack's weight-gradient stream, and reported alternately (text, audio, text, ...) once both are enqueued.
    struct StackRange { int64_t lo, hi; cudaEvent_t ev; };
    std::vector<StackRange> stack_ranges[2];
    int collecting = -1;
    cudaStream_t collect_stream = nullptr;
    size_t stack_ev_used = 0;
    void collect_begin(int modality, cudaStream_t wst) { collecting = modality; collect_stream = wst; }
    void collect_end() { collecting = -1; }
    int report_stacks() {
        cudaStream_t ready = p->use_streams ? p->ready : st;
        const size_t n = std::max(stack_ranges[0].size(), stack_ranges[1].size());
        for (size_t i = 0; i < n; ++i)
            for (int s = 1; s >= 0; --s) {
                if (i >= stack_ranges[s].size()) continue;
                const StackRange& r = stack_ranges[s][i];
                if (ready != st || p->use_streams) M2F_CUDA(cudaStreamWaitEvent(ready, r.ev, 0));
                M2F_TRY(report(r.lo, r.hi, ready));
            }
        stack_ranges[0].clear();
        stack_ranges[1].clear();
        return M2F_OK;
    }
    // eager: tell the caller now; capture: leave an external event-record node behind, the caller is told after the launch
    std::vector<std::pair<int64_t, int64_t>>* marks = nullptr;
    bool ready_in_capture = false;
    bool xchg = false;       // exchange the range here (kernel on `ready`, captured with the graph) instead of telling the caller
    bool capturing = false;
    int report(int64_t lo, int64_t hi_, cudaStream_t ready, bool last = false) {
        if (xchg) {  // one sequence on `ready`: only the last call needs the closing handshake
            M2F_TRY(m2f_grad_allreduce_peer(&p->xchg_group, lo, hi_, p->xchg_scale, p->xchg_ctas,
                                            (last ? 0 : M2F_XCHG_NO_EXIT_BARRIER) | (p->xchg_group.wire_bf16 ? M2F_XCHG_WIRE_BF16 : 0), ready));
            ++p->xchg_ranges;
            if (ready != st) ready_in_capture = true;
            if (tail_adam) {
                while (p->xchg_evs.size() < xranges.size() + 1) {
                    cudaEvent_t e;
                    M2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    p->xchg_evs.push_back(e);
                }
                M2F_CUDA(cudaEventRecord(p->xchg_evs[xranges.size()], ready));
                xranges.emplace_back(lo, hi_);
            }
            return M2F_OK;
        }
        if (!marks) {
            cb(cb_user, lo, hi_, ready);
            return M2F_OK;
        }
        while (p->bucket_evs.size() <= marks->size()) {
            cudaEvent_t e;
            M2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            p->bucket_evs.push_back(e);
        }
        M2F_CUDA(cudaEventRecordWithFlags(p->bucket_evs[marks->size()], ready, cudaEventRecordExternal));
        marks->emplace_back(lo, hi_);
        if (ready != st) ready_in_capture = true;
        return M2F_OK;
    }

    // ---- optimizer behind the exchange.  With an exchange attached, what is left behind the backward's last kernel is the
    // backlog of the exchange stream (the last few 21 MB ranges, ~0.3 ms at C2), and after it optimizer.step() streams the whole
    // arena (~0.5 ms).  Here the update of range k is enqueued on the caller's stream -- behind ALL of the backward's compute, so
    // nothing reads the weights any more and no GEMM competes for L2 -- gated only by the event behind exchange call k + 1: a
    // later call's entry barrier proves that every rank has completed call k (stores included), so range k is final here while
    // the calls after it are still in flight; the step's last call closes with its own handshake.  Same arithmetic, bit for
    // bit, as m2f_model_adam_step after the backward.
    bool tail_adam = false;
    std::vector<std::pair<int64_t, int64_t>> xranges;
    int adam_tail() {
        const size_t n = xranges.size();
        for (size_t k = 0; k < n; ++k) {
            cudaEvent_t gate = p->xchg_evs[k + 1 < n ? k + 1 : n - 1];
            if (p->use_streams) M2F_CUDA(cudaStreamWaitEvent(st, gate, 0));
            M2F_TRY(adam_range(xranges[k].first, xranges[k].second, st));
        }
        return M2F_OK;
    }
    int adam_range(int64_t lo, int64_t hi_, cudaStream_t ast) {
        m2f_model* md = p->m;
        const AdamArgs* hp = reinterpret_cast<const AdamArgs*>(p->opt_hp);
        float* prm_w = const_cast<float*>(params);
        if (lo >= md->off_1d) {  // biases / LayerNorm vectors, then their packed fp32 copies in the shadow
            const int64_t n4 = (hi_ - lo) / 4;
            int64_t g = (n4 + 255) / 256;
            if (g > 148 * 8) g = 148 * 8;
            adam_range_kernel<<<(unsigned)g, 256, 0, ast>>>(reinterpret_cast<float4*>(prm_w + lo), reinterpret_cast<const float4*>(grads + lo),
                                                           reinterpret_cast<float4*>(p->opt_m + lo), reinterpret_cast<float4*>(p->opt_v + lo), n4, hp);
            M2F_LAUNCH_CHECK();
            ShadowBatch b;
            b.n = 0;
            for (const SEntry& e : md->sentries) {
                if (e.planes != 0) continue;
                b.e[b.n++] = e;
                if (b.n == SHADOW_CHUNK) {
                    shadow_refresh_kernel<<<dim3(4, b.n), 256, 0, ast>>>(params, shadow, b);
                    M2F_LAUNCH_CHECK();
                    b.n = 0;
                }
            }
            if (b.n) {
                shadow_refresh_kernel<<<dim3(4, b.n), 256, 0, ast>>>(params, shadow, b);
                M2F_LAUNCH_CHECK();
            }
            return M2F_OK;
        }
        ShadowBatch b;
        b.n = 0;
        const AdamArgs unused{};
        auto flush = [&]() -> int {
            if (b.n == 0) return M2F_OK;
            adam_shadow_kernel<<<dim3(96, b.n), 256, 0, ast>>>(prm_w, grads, p->opt_m, p->opt_v, shadow, b, unused, hp);
            M2F_LAUNCH_CHECK();
            b.n = 0;
            return M2F_OK;
        };
        for (const SEntry& e : md->sentries) {
            if (e.planes == 0 || e.src_off < lo || e.src_off >= hi_) continue;
            b.e[b.n++] = e;
            if (b.n == SHADOW_CHUNK) M2F_TRY(flush());
        }
        return flush();
    }

    m2f_planes_t pv(const Buf& b) const {
        m2f_planes_t t;
        t.ptr = (b.base == WS ? ws : shadow) + b.off;
        t.rows = b.rows; t.cols = b.cols; t.ld = b.ld; t.plane_stride = b.stride; t.planes = b.planes; t._pad = 0;
        return t;
    }
    // Backward operand view: the leading planes_bwd planes of a saved activation / weight matrix (scratch buffers of the
    // backward are allocated with planes_bwd planes, for them this is the whole matrix).
    m2f_planes_t pvb(const Buf& b) const {
        m2f_planes_t t = pv(b);
        const int pb = p->m->cfg.planes_bwd > 0 ? p->m->cfg.planes_bwd : p->m->cfg.planes;
        if (t.planes > pb) t.planes = pb;
        return t;
    }
    static m2f_planes_t cols(m2f_planes_t t, int64_t c0, int64_t n) {
        t.ptr = reinterpret_cast<uint8_t*>(t.ptr) + c0 * 2;
        t.cols = n;
        return t;
    }
    float* f32(const Buf& b) const { return reinterpret_cast<float*>((b.base == WS ? ws : shadow) + b.off); }
    const float* prm(int64_t off) const { return params + off; }
    float* grd(int64_t off) const { return grads + off; }

    int tmap(const m2f_planes_t& t, int box_rows, const CUtensorMap** out, bool sw64 = false) {
        TmapKey k{t.ptr, t.rows, t.cols, t.ld, t.plane_stride, t.planes, box_rows | (sw64 ? 0x10000 : 0)};
        auto it = p->tmaps.find(k);
        if (it == p->tmaps.end()) {
            CUtensorMap m;
            if (sw64) M2F_TRY(make_planes_tmap_sw64(&m, &t, box_rows));
            else M2F_TRY(make_planes_tmap(&m, &t, box_rows));
            it = p->tmaps.emplace(k, m).first;
        }
        *out = &it->second;
        return M2F_OK;
    }

    int gemm(const m2f_planes_t& A, bool a_mn, const m2f_planes_t& B, bool b_mn, int64_t M, int64_t N, const m2f_epilogue_t& e,
             const m2f_kseg_t* segs = nullptr, int nsegs = 0) {
        m2f_kseg_t one;
        if (!segs) {
            const int64_t ka = a_mn ? A.rows : A.cols, kb = b_mn ? B.rows : B.cols;
            M2F_REQUIRE(ka == kb, M2F_ERR_ARG, "plan gemm: contraction extents differ (%lld vs %lld)", (long long)ka, (long long)kb);
            one = {0, 0, (int32_t)ka};
            segs = &one;
            nsegs = 1;
        }
        GemmParams gp;
        M2F_TRY(gemm_fill_params(gp, &A, a_mn, &B, b_mn, segs, nsegs, M, N, &e));
        {   // weights live in the shadow buffer, which only the refresh after the optimizer step writes
            const uint8_t* bp = reinterpret_cast<const uint8_t*>(B.ptr);
            gp.b_static = (bp >= shadow && bp < shadow + p->m->shadow_bytes) ? 1 : 0;
        }
        int64_t ktot = 0;
        for (int i = 0; i < nsegs; ++i) ktot += segs[i].klen;
        int bn = gemm_block_n(M, N, ktot);
        // Serial sections (fusion layers, classifier: nothing else runs beside the chain): a 768-wide GEMM at M = 2112 is
        // 9 x 3 pair tiles = 54 CTAs, a third of the SMs.  Half-width tiles double the CTAs and halve each CTA's main loop.
        if (narrow && (bn & GEMM_TILE_CTA2) && (bn & 0xfff) == 256 && N > 128 &&
            ((M + 2 * BLOCK_M - 1) / (2 * BLOCK_M)) * 2 * ((N + 255) / 256) <= 74)
            bn = 128 | GEMM_TILE_CTA2;
        // Forward GEMMs (both operands K-major) whose N is a multiple of 192 (768: out-proj, FFN2, projections; 2304: packed QKV):
        // 256 x 192 pair tiles give 4 / 12 column tiles instead of 3 / 9 -- at M = 2112 the two stacks' 768-wide GEMMs together
        // are 144 CTAs (one wave of the 148 SMs, each with 3/4 of the main loop) instead of 108, and the QKV pair 432 CTAs of 3/4
        // (three waves, the last one full) instead of 324 (three waves, the last one a fifth full).  M2F_GEMM_BN192: 0 off,
        // 1 (default) inside the modality stacks, 2 also instead of the half-width tiles of the serial sections.
        // The same tiles exist for the backward (M2F_GEMM_BN192_BWD: bit 1 data gradients -- B = W read MN-major, 96 columns per CTA
        // through the 64-byte swizzle --, bit 2 weight gradients) and are OFF by default: measured 3.357 (off) / 3.366 (dgrad) /
        // 3.418 ms (wgrad) per step -- the single-product backward GEMMs are bound by their fixed costs, not their main loops, and
        // more weight-gradient CTAs take SMs from the data-gradient chain (profiles/r2_ab_tiles.txt).
        static const int bn192 = [] { const char* e = std::getenv("M2F_GEMM_BN192"); return e ? std::atoi(e) : 1; }();
        static const int bn192b = [] { const char* e = std::getenv("M2F_GEMM_BN192_BWD"); return e ? std::atoi(e) : 0; }();
        const bool fwd_like = !a_mn && !b_mn, dgrad_like = !a_mn && b_mn, wgrad_like = a_mn && b_mn;
        if ((bn & GEMM_TILE_CTA2) && !(bn & GEMM_TILE_PERSIST) && N % 192 == 0 && N >= 384 &&
            ((bn192 && fwd_like && (((bn & 0xfff) == 256 && !narrow) || (bn192 == 2 && narrow)) && !(bn192 == 3 && N > 768) && !(bn192 == 4 && N <= 768)) ||
             (bn192b && (bn & 0xfff) == 256 && ((dgrad_like && (bn192b & 1)) || (wgrad_like && (bn192b & 2))))))
            bn = 192 | GEMM_TILE_CTA2;
        EpiKey ek;
        std::memset(&ek, 0, sizeof(ek));  // padding bytes take part in the comparison
        ek.out_f32 = gp.out_f32; ek.out_pl = gp.out_pl; ek.gate = gp.gate; ek.resid = gp.resid;
        ek.out_f32_ld = gp.out_f32_ld; ek.out_pl_ld = gp.out_pl_ld; ek.out_pl_stride = gp.out_pl_stride; ek.gate_ld = gp.gate_ld;
        ek.resid_ld = gp.resid_ld; ek.M = gp.M; ek.N = gp.N; ek.out_pl_n = gp.out_pl_n;
        auto eit = p->epimaps.find(ek);
        if (eit == p->epimaps.end()) {
            EpiEntry ent;
            std::memset(&ent, 0, sizeof(ent));
            M2F_TRY(gemm_make_epi_maps(gp, &ent.maps));
            ent.epi_tma = gp.epi_tma;
            eit = p->epimaps.emplace(ek, ent).first;
        }
        gp.epi_tma = eit->second.epi_tma;
        const CUtensorMap *ta, *tb;
        M2F_TRY(tmap(A, a_mn ? BLOCK_K : BLOCK_M, &ta));
        M2F_TRY(tmap(B, b_mn ? BLOCK_K : gemm_b_box_rows(bn), &tb, b_mn && (bn & 0xfff) == 192));
        return gemm_launch(*ta, *tb, eit->second.maps, gp, bn, a_mn, b_mn, stream());
    }
};

struct Epi {
    m2f_epilogue_t e;
    Epi() { memset(&e, 0, sizeof(e)); e.gate_scale = 1.f; }
    Epi& bias(const float* b) { e.bias = b; return *this; }
    Epi& relu() { e.relu = 1; return *this; }
    Epi& gate(const m2f_planes_t& g, float scale) { e.gate = g.ptr; e.gate_ld = g.ld; e.gate_scale = scale; return *this; }
    Epi& drop(float p, uint64_t seed, uint64_t step, uint32_t site, int64_t ld) {
        e.drop_p = p; e.seed = seed; e.step = step; e.site = site; e.drop_ld = ld; return *this;
    }
    Epi& resid(const float* r, int64_t ld) { e.resid = r; e.resid_ld = ld; return *this; }
    Epi& f32(float* o, int64_t ld, bool acc = false) { e.out_f32 = o; e.out_f32_ld = ld; e.accumulate = acc ? 1 : 0; return *this; }
    Epi& pl(const m2f_planes_t& o) { e.out_planes = o; return *this; }
    Epi& colsum(float* cs) { e.colsum = cs; return *this; }
    operator const m2f_epilogue_t&() const { return e; }
};

static inline float inv_keep(float p) { return 1.f / (1.f - p); }

// Timing aid (results become garbage): M2F_DEBUG_SKIP=<prefix>[,<prefix>...] leaves out every op of the step whose call text starts
// with one of the prefixes (e.g. "m2f_ln_fwd,m2f_attn_bwd") -- the step time without them bounds what a faster kernel could win.
static bool debug_skip(const char* expr) {
    static const std::vector<std::string> pats = [] {
        std::vector<std::string> v;
        const char* e = std::getenv("M2F_DEBUG_SKIP");
        std::string cur;
        for (const char* q = e ? e : ""; ; ++q) {
            if (*q == ',' || *q == 0) {
                if (!cur.empty()) v.push_back(cur);
                cur.clear();
                if (*q == 0) break;
            } else {
                cur.push_back(*q);
            }
        }
        return v;
    }();
    for (const std::string& pat : pats)
        if (std::strncmp(expr, pat.c_str(), pat.size()) == 0) return true;
    return false;
}

// an op of the step: M2F_TRY + (debug) a timeline stamp behind it on the same stream
#define OP(expr)                               \
    do {                                       \
        if (!debug_skip(#expr)) M2F_TRY(expr); \
        c.stamp(#expr);                        \
    } while (0)

// ------------------------------------------------------------------------------------------ forward
static int forward_modality(Ctx& c, Modality& mo, const float* input, const uint8_t* key_pad, const m2f_planes_t& proj_dst) {
    m2f_plan* p = c.p;
    const int64_t M = p->M;
    const int D = mo.D;
    const float pd = p->p_used;
    const uint64_t seed = p->seed, step = p->step;
    m2f_planes_t in_pl = c.pv(mo.in_pl);
    OP(m2f_split_planes(input, D, &in_pl, c.stream()));
    const float* x_f32 = input;
    int64_t x_ld = D;
    m2f_planes_t x_pl = in_pl;
    for (int t = 0; t < mo.T; ++t) {
        const float* cur = x_f32;
        int64_t cur_ld = x_ld;
        m2f_planes_t cur_pl = x_pl;
        for (int l = 0; l < mo.L; ++l) {
            EncLayer& e = mo.layers[(size_t)t * mo.L + l];
            m2f_planes_t qkv = c.pv(e.qkv_pl), o = c.pv(e.o_pl), x1p = c.pv(e.x1_pl), h = c.pv(e.h_pl), x2p = c.pv(e.x2_pl);
            // packed QKV projection (torch/nn/functional.py:5833-5846)
            OP(c.gemm(cur_pl, false, c.pv(e.qkv.w), false, M, 3 * D, Epi().bias(c.prm(e.qkv.b_off)).pl(qkv)));
            m2f_planes_t q = Ctx::cols(qkv, 0, D), k = Ctx::cols(qkv, D, D), v = Ctx::cols(qkv, 2 * D, D);
            OP(m2f_attn_fwd(&q, &k, &v, key_pad, p->B, p->S, mo.H, pd, seed, step, M2F_SITE_ENC(mo.idx, t, l, 0), &o, c.stream()));
            // x = norm1(x + dropout1(out_proj(attn)))   (transformer.py:951-977)
            OP(c.gemm(o, false, c.pv(e.wo.w), false, M, D,
                           Epi().bias(c.prm(e.wo.b_off)).drop(pd, seed, step, M2F_SITE_ENC(mo.idx, t, l, 1), D).resid(cur, cur_ld).f32(c.f32(e.y1), e.y1.ld)));
            OP(m2f_ln_fwd(c.f32(e.y1), e.y1.ld, c.prm(e.g1), c.prm(e.be1), nullptr, 0, M, D, c.f32(e.x1), e.x1.ld, &x1p, 0.f, 0, 0, 0,
                               c.f32(e.st1), c.stream()));
            // x = norm2(x + dropout2(linear2(dropout(relu(linear1(x))))))   (transformer.py:980-982)
            OP(c.gemm(x1p, false, c.pv(e.w1.w), false, M, e.w1.out,
                           Epi().bias(c.prm(e.w1.b_off)).relu().drop(pd, seed, step, M2F_SITE_ENC(mo.idx, t, l, 2), e.w1.out).pl(h)));
            OP(c.gemm(h, false, c.pv(e.w2.w), false, M, D,
                           Epi().bias(c.prm(e.w2.b_off)).drop(pd, seed, step, M2F_SITE_ENC(mo.idx, t, l, 3), D).resid(c.f32(e.x1), e.x1.ld).f32(c.f32(e.y2), e.y2.ld)));
            OP(m2f_ln_fwd(c.f32(e.y2), e.y2.ld, c.prm(e.g2), c.prm(e.be2), nullptr, 0, M, D, c.f32(e.x2), e.x2.ld, &x2p, 0.f, 0, 0, 0,
                               c.f32(e.st2), c.stream()));
            cur = c.f32(e.x2);
            cur_ld = e.x2.ld;
            cur_pl = x2p;
        }
        // x = x + norm(encoder(x))  (src/model.py:107,119; the norm is shared across t); the last one
        // also applies the dropout in front of the projection (src/model.py:111,123) to its operand copy
        const bool last = (t == mo.T - 1);
        m2f_planes_t outp = c.pv(mo.xout_pl[t]);
        OP(m2f_ln_fwd(cur, cur_ld, c.prm(mo.gf), c.prm(mo.bef), x_f32, x_ld, M, D, last ? nullptr : c.f32(mo.xout[t]), mo.xout[t].ld, &outp,
                           last ? pd : 0.f, seed, step, M2F_SITE_MOD(mo.idx, 0), c.f32(mo.stf[t]), c.stream()));
        x_f32 = c.f32(mo.xout[t]);
        x_ld = mo.xout[t].ld;
        x_pl = outp;
    }
    // projection to the fusion width + dropout (src/model.py:112-113,124-125)
    OP(c.gemm(x_pl, false, c.pv(mo.proj.w), false, M, mo.proj.out,
                   Epi().bias(c.prm(mo.proj.b_off)).drop(pd, seed, step, M2F_SITE_MOD(mo.idx, 1), mo.proj.out).pl(proj_dst)));
    return M2F_OK;
}

enum { PH_HEAD = 1, PH_BODY = 2, PH_TAIL = 4, PH_ALL = 7 };

// M2F_FUSION_FORK (default 1): inside the serial fusion section, GEMMs that do not depend on each other run on an idle side
// stream beside the chain -- forward: the stacked K projections beside relu + the first layer's q|v projection; backward: the
// text half of d u (needed only by the layer's last GEMM) beside out-proj data gradient + attention backward.  0: one chain.
static bool fusion_fork() {
    static const bool on = [] { const char* e = std::getenv("M2F_FUSION_FORK"); return !(e && e[0] == '0'); }();
    return on;
}

// M2F_NARROW: 0 = full-width tiles everywhere; 1 (default) = half-width tiles for the under-filled GEMMs of the serial sections
// (fusion layers, classifier; measured 3.85 -> 3.74 ms per step at C2, profiles/r2_ab_tiles.txt); 2 = also inside the two
// modality stacks (measured slower: 4.05 vs 4.00 in the same run); 3 = also for the weight gradients (slower still).
static int narrow_level() {
    static const int lvl = [] { const char* e = std::getenv("M2F_NARROW"); return e ? std::atoi(e) : 1; }();
    return lvl;
}
static bool narrow_enabled() { return narrow_level() >= 1; }

// PH_BODY: everything up to the classifier's last hidden layer (capturable: touches only plan-owned memory when text /
// audio point at the staging copies); PH_TAIL: the logits GEMM (writes the caller's buffer).
static int forward_all(Ctx& c, const float* text, const float* audio, const uint8_t* key_pad, float* logits, int phases) {
    m2f_plan* p = c.p;
    m2f_model* md = p->m;
    const int64_t M = p->M;
    const int Df = md->Df;
    const float pd = p->p_used;
    const uint64_t seed = p->seed, step = p->step;
    m2f_planes_t zc = c.pv(p->zc_pl);
    const bool fam = !p->fam.empty();
    // the two modality stacks are independent until the first fusion layer: audio runs on its own stream
    const bool fork = md->both && p->use_streams;
    const bool body = phases & PH_BODY;
    if (body) c.stamp("begin");
    c.narrow = narrow_level() >= 2;
    if (body && p->mod[0].enabled) {
        if (fork) {
            M2F_TRY(c.dep(c.st, c.side(0)));
            c.cur = c.side(0);
        }
        M2F_TRY(forward_modality(c, p->mod[0], audio, key_pad, md->both ? Ctx::cols(zc, 0, Df) : zc));
        c.cur = nullptr;
    }
    if (body && p->mod[1].enabled) {
        m2f_planes_t dst = fam ? c.pv(p->t0_pl) : (md->both ? Ctx::cols(zc, Df, Df) : zc);
        M2F_TRY(forward_modality(c, p->mod[1], text, key_pad, dst));
    }
    if (body && fork) M2F_TRY(c.dep(c.side(0), c.st));
    c.narrow = narrow_enabled();
    if (body && fam) {
        const int F = (int)p->fam.size();
        m2f_planes_t a = Ctx::cols(zc, 0, Df);
        m2f_planes_t kall = c.pv(p->kall_pl);
        // K projections of every fusion layer read the same audio features: one GEMM
        const bool kfork = body && fork && fusion_fork();  // the audio chain's stream is idle from the join on
        if (kfork) {
            M2F_TRY(c.dep(c.st, c.side(0)));
            c.cur = c.side(0);
        }
        OP(c.gemm(a, false, c.pv(md->wk_all.w), false, M, (int64_t)F * Df, Epi().bias(c.f32(md->bk_all)).pl(kall)));
        c.cur = nullptr;
        m2f_planes_t t_cur = c.pv(p->t0_pl);
        {   // relu(text) half of the first concat operand (src/model.py:16-17); later layers' text is already >= 0
            m2f_planes_t u0r = Ctx::cols(c.pv(p->fam[0].u_pl), Df, Df);
            OP(m2f_relu_planes(&t_cur, &u0r, c.stream()));
        }
        for (int f = 0; f < F; ++f) {
            FusionLayer& fl = p->fam[f];
            m2f_planes_t qv = c.pv(fl.qv_pl), xo = c.pv(fl.xo_pl), u = c.pv(fl.u_pl);
            OP(c.gemm(t_cur, false, c.pv(fl.qv.w), false, M, 2 * Df, Epi().bias(c.f32(fl.bqv)).pl(qv)));
            // query = text, key = audio, value = text  (src/model.py:14)
            m2f_planes_t q = Ctx::cols(qv, 0, Df), v = Ctx::cols(qv, Df, Df), k = Ctx::cols(kall, (int64_t)f * Df, Df);
            if (f == 0 && kfork) M2F_TRY(c.dep(c.side(0), c.st));  // the K projections are needed from here on
            OP(m2f_attn_fwd(&q, &k, &v, key_pad, p->B, p->S, md->cfg.fam_heads, pd, seed, step, M2F_SITE_FAM(f, 0), &xo, c.stream()));
            OP(c.gemm(xo, false, c.pv(fl.wo.w), false, M, Df, Epi().bias(c.prm(fl.wo.b_off)).relu().pl(Ctx::cols(u, 0, Df))));
            m2f_planes_t t_next = (f < F - 1) ? Ctx::cols(c.pv(p->fam[f + 1].u_pl), Df, Df) : Ctx::cols(zc, Df, Df);
            // relu(linear(relu(cat(x, text)))) then the model-level dropout (src/model.py:16-19,131)
            OP(c.gemm(u, false, c.pv(fl.lin.w), false, M, Df,
                           Epi().bias(c.prm(fl.lin.b_off)).relu().drop(pd, seed, step, M2F_SITE_FAM(f, 1), Df).pl(t_next)));
            t_cur = t_next;
        }
    }
    // classifier (src/model.py:89-100,143)
    const int n_lin = (int)md->cls.size();
    m2f_planes_t in = zc;
    for (int j = 0; j < n_lin; ++j) {
        Lin& L = md->cls[j];
        if (j < n_lin - 1) {
            m2f_planes_t hj = c.pv(p->cls_h[j]);
            Epi e;
            e.bias(c.prm(L.b_off)).relu().pl(hj);
            if (j == n_lin - 2) e.drop(pd, seed, step, M2F_SITE_CLS, L.out);
            if (body) OP(c.gemm(in, false, c.pv(L.w), false, M, L.out, e));
            in = hj;
        } else if (phases & PH_TAIL) {
            OP(c.gemm(in, false, c.pv(L.w), false, M, L.out, Epi().bias(c.prm(L.b_off)).f32(logits, L.out)));
        }
    }
    return M2F_OK;
}

// ------------------------------------------------------------------------------------------ backward
// wgrad + bias grad of one Linear: dW[out,in] = dY^T X, db = colsum(dY)
// Off the critical path: enqueued on the weight-gradient stream `wst`, ordered after what the current stream holds so far.
static int lin_wgrad(Ctx& c, cudaStream_t wst, const m2f_planes_t& dY, const m2f_planes_t& X, float* dW, int64_t dW_ld, float* db, int64_t out,
                     int64_t in) {
    cudaStream_t chain = c.stream();
    M2F_TRY(c.dep(chain, wst));
    c.cur = wst;
    const bool narrow = c.narrow;
    c.narrow = narrow_level() >= 3;  // weight gradients run beside the chain: full-width tiles
    int rc = debug_skip("wgrad") ? M2F_OK : c.gemm(dY, true, X, true, out, in, Epi().f32(dW, dW_ld));
    c.narrow = narrow;
    if (rc == M2F_OK && db) rc = m2f_colsum_planes(&dY, db, wst);
    c.stamp("wgrad");
    c.cur = chain == c.st ? nullptr : chain;
    return rc;
}

// M2F_FINE_TAIL=1: the layer computed last reports its four weight gradients one by one (shorter exchange after the end of
// the backward, three more exchange calls per stack)
static bool fine_tail() {
    static const bool on = [] { const char* e = std::getenv("M2F_FINE_TAIL"); return e && e[0] == '1'; }();
    return on;
}

// One modality stack, last layer first.  The data-gradient chain runs on the current stream; every weight gradient goes
// to `wst` (they are off the critical path).  The operand buffers the weight-gradient stream reads alternate between two
// sets by layer parity; a set is reused only after the weight gradients of its previous user (two layers earlier) are done.
static int backward_modality(Ctx& c, Modality& mo, const uint8_t* key_pad, const m2f_planes_t& dproj, cudaStream_t wst) {
    m2f_plan* p = c.p;
    const int64_t M = p->M;
    const int D = mo.D;
    const float pd = p->p_used;
    const uint64_t seed = p->seed, step = p->step;
    m2f_plan::StackScratch& sc = p->ss[mo.idx];
    float* G = c.f32(sc.gG);
    float* A = c.f32(sc.gA);
    float* Bq = c.f32(sc.gB);
    const int64_t gld = sc.gG.ld;
    // projection: dW, db, and g = dropmask0(dproj . Wp)
    m2f_planes_t xp = c.pvb(mo.xout_pl[mo.T - 1]);
    M2F_TRY(lin_wgrad(c, wst, dproj, xp, c.grd(mo.proj.w_off), mo.proj.in, nullptr, mo.proj.out, mo.proj.in));  // db: by dproj's producer
    M2F_TRY(c.done_down_to(mo.proj.w_off));
    OP(c.gemm(dproj, false, c.pvb(mo.proj.w), true, M, D, Epi().drop(pd, seed, step, M2F_SITE_MOD(mo.idx, 0), D).f32(G, gld)));
    int nlayer = 0;
    for (int t = mo.T - 1; t >= 0; --t) {
        const EncLayer& lastl = mo.layers[(size_t)t * mo.L + mo.L - 1];
        // x_out = x_in + LN_f(x2_last): G is d x_out; A <- d x2_last; G stays as the skip gradient
        OP(m2f_ln_bwd(G, gld, c.f32(lastl.x2), lastl.x2.ld, c.prm(mo.gf), c.f32(mo.stf[t]), M, D, A, gld, nullptr, 0.f, 0, 0, 0,
                           c.grd(mo.gf), c.grd(mo.bef), nullptr, c.stream()));
        for (int l = mo.L - 1; l >= 0; --l, ++nlayer) {
            EncLayer& e = mo.layers[(size_t)t * mo.L + l];
            const bool first = (t == 0 && l == 0);
            const int par = nlayer & 1;
            if (nlayer >= 2 && wst != c.stream()) M2F_CUDA(cudaStreamWaitEvent(c.stream(), p->wg_ev[mo.idx][par], 0));
            m2f_planes_t dx2 = c.pvb(sc.dx2[par]), dx1 = c.pvb(sc.dx1[par]), dhp = c.pvb(sc.dh[par]), dop = c.pvb(sc.do_[par]), dqkvp = c.pvb(sc.dqkv[par]);
            m2f_planes_t in_pl = (l > 0) ? c.pvb(mo.layers[(size_t)t * mo.L + l - 1].x2_pl) : (t > 0 ? c.pvb(mo.xout_pl[t - 1]) : c.pvb(mo.in_pl));
            m2f_planes_t h = c.pvb(e.h_pl), x1p = c.pvb(e.x1_pl), o = c.pvb(e.o_pl), qkv = c.pvb(e.qkv_pl);
            // LN2: A = d x2 -> Bq = d y2 (fp32), dx2 = dropmask3(d y2), db2 += colsum
            OP(m2f_ln_bwd(A, gld, c.f32(e.y2), e.y2.ld, c.prm(e.g2), c.f32(e.st2), M, D, Bq, gld, &dx2, pd, seed, step,
                               M2F_SITE_ENC(mo.idx, t, l, 3), c.grd(e.g2), c.grd(e.be2), c.grd(e.w2.b_off), c.stream()));
            M2F_TRY(lin_wgrad(c, wst, dx2, h, c.grd(e.w2.w_off), e.w2.in, nullptr, e.w2.out, e.w2.in));
            // the layer computed last: hand its weights over one by one, so that only the final one is left behind the backward
            if (first && fine_tail()) M2F_TRY(c.done_down_to(e.w2.w_off, true));
            // dh = (d y2 . W2) * [h > 0] / (1-p)   (relu + dropout backward from the saved output)
            OP(c.gemm(dx2, false, c.pvb(e.w2.w), true, M, e.w2.in, Epi().gate(h, inv_keep(pd)).pl(dhp).colsum(c.grd(e.w1.b_off))));
            M2F_TRY(lin_wgrad(c, wst, dhp, x1p, c.grd(e.w1.w_off), e.w1.in, nullptr, e.w1.out, e.w1.in));
            if (first && fine_tail()) M2F_TRY(c.done_down_to(e.w1.w_off, true));
            // A = d x1 = dh . W1 + d y2
            OP(c.gemm(dhp, false, c.pvb(e.w1.w), true, M, D, Epi().resid(Bq, gld).f32(A, gld)));
            // LN1: A -> Bq = d y1, dx1 = dropmask1(d y1), dbo += colsum
            OP(m2f_ln_bwd(A, gld, c.f32(e.y1), e.y1.ld, c.prm(e.g1), c.f32(e.st1), M, D, Bq, gld, &dx1, pd, seed, step,
                               M2F_SITE_ENC(mo.idx, t, l, 1), c.grd(e.g1), c.grd(e.be1), c.grd(e.wo.b_off), c.stream()));
            M2F_TRY(lin_wgrad(c, wst, dx1, o, c.grd(e.wo.w_off), e.wo.in, nullptr, e.wo.out, e.wo.in));
            if (first && fine_tail()) M2F_TRY(c.done_down_to(e.wo.w_off, true));
            OP(c.gemm(dx1, false, c.pvb(e.wo.w), true, M, D, Epi().pl(dop)));
            m2f_planes_t q = Ctx::cols(qkv, 0, D), k = Ctx::cols(qkv, D, D), v = Ctx::cols(qkv, 2 * D, D);
            m2f_planes_t dq = Ctx::cols(dqkvp, 0, D), dk = Ctx::cols(dqkvp, D, D), dv = Ctx::cols(dqkvp, 2 * D, D);
            float* dbqkv = c.grd(e.qkv.b_off);  // in_proj_bias gradient = column sums of [dq | dk | dv], folded into the attention backward
            OP(m2f_attn_bwd(&q, &k, &v, &dop, key_pad, p->B, p->S, mo.H, pd, seed, step, M2F_SITE_ENC(mo.idx, t, l, 0), &dq, &dk, &dv,
                                 dbqkv, dbqkv + D, dbqkv + 2 * D, c.stream()));
            M2F_TRY(lin_wgrad(c, wst, dqkvp, in_pl, c.grd(e.qkv.w_off), e.qkv.in, nullptr, e.qkv.out, e.qkv.in));
            if (wst != c.stream()) M2F_CUDA(cudaEventRecord(p->wg_ev[mo.idx][par], wst));
            M2F_TRY(c.done_down_to(e.qkv.w_off));
            if (!first) {
                // gradient w.r.t. the layer input = dqkv . Wqkv + d y1 ; at l == 0 it also joins the skip gradient in G
                if (l > 0) OP(c.gemm(dqkvp, false, c.pvb(e.qkv.w), true, M, D, Epi().resid(Bq, gld).f32(A, gld)));
                else       OP(c.gemm(dqkvp, false, c.pvb(e.qkv.w), true, M, D, Epi().resid(Bq, gld).f32(G, gld, true)));
            }
        }
    }
    return M2F_OK;
}

// PH_HEAD: dlogits -> planes (reads the caller's buffer); PH_BODY: the rest (capturable)
static int backward_all(Ctx& c, const uint8_t* key_pad, const float* dlogits, int phases) {
    m2f_plan* p = c.p;
    m2f_model* md = p->m;
    const int64_t M = p->M;
    const int Df = md->Df;
    const float pd = p->p_used;
    const uint64_t seed = p->seed, step = p->step;
    const bool fam = !p->fam.empty();
    m2f_planes_t dl = c.pvb(p->dl_pl);
    if (phases & PH_HEAD) {
        if (dlogits) {
            m2f_planes_t dlc = dl;
            dlc.cols = md->cfg.cls_out;
            OP(m2f_split_planes(dlogits, md->cfg.cls_out, &dlc, c.stream()));
        } else {
            M2F_REQUIRE(p->have_loss_grad, M2F_ERR_ARG, "plan_backward: dlogits is NULL but m2f_plan_loss was not called after the forward");
        }
    }
    if (!(phases & PH_BODY)) return M2F_OK;
    // 1-D gradients are accumulated with atomics: zero that region of the arena
    M2F_CUDA(cudaMemsetAsync(c.grads + md->off_1d, 0, (size_t)(md->arena_floats - md->off_1d) * 4, c.st));
    dl.cols = md->cfg.cls_out;
    m2f_planes_t zc = c.pvb(p->zc_pl);
    // ---- classifier
    const int n_lin = (int)md->cls.size();
    m2f_planes_t d_cur = dl;
    m2f_planes_t dproj[2];
    dproj[0] = c.pvb(p->dproj_pl);
    dproj[1] = c.pvb(p->dprojt_pl);
    m2f_planes_t dz = c.pvb(p->dz[0]);
    cudaStream_t ws1 = c.side(1);  // weight gradients of the classifier, the fusion layers and the text stack
    c.narrow = narrow_enabled();
    for (int j = n_lin - 1; j >= 0; --j) {
        Lin& L = md->cls[j];
        m2f_planes_t in = (j == 0) ? zc : c.pvb(p->cls_h[j - 1]);
        // bias gradients ride in the epilogue of the GEMM that produces dY; only dlogits comes from elsewhere
        M2F_TRY(lin_wgrad(c, ws1, d_cur, in, c.grd(L.w_off), L.in, j == n_lin - 1 ? c.grd(L.b_off) : nullptr, L.out, L.in));
        if (j > 0) {
            if (j <= n_lin - 3) M2F_TRY(c.dep(ws1, c.st));  // the buffer written next was an operand of the weight gradient two layers up
            m2f_planes_t dn = c.pvb((j & 1) ? p->dcls_a : p->dcls_b);
            const float sc = (j == n_lin - 1) ? inv_keep(pd) : 1.f;  // only the last hidden layer has dropout
            OP(c.gemm(d_cur, false, c.pvb(L.w), true, M, L.in, Epi().gate(in, sc).pl(dn).colsum(c.grd(md->cls[j - 1].b_off))));
            d_cur = dn;
        } else if (md->both) {
            m2f_planes_t w = c.pvb(L.w);
            // audio half of the concat: gradient w.r.t. the projection's pre-dropout output (mask regenerated)
            if (fam) {
                OP(c.gemm(d_cur, false, Ctx::cols(w, 0, Df), true, M, Df,
                               Epi().drop(pd, seed, step, M2F_SITE_MOD(0, 1), Df).f32(c.f32(p->ga_f32), p->ga_f32.ld)));
                // text half: gradient w.r.t. the last fusion linear's pre-activation
                OP(c.gemm(d_cur, false, Ctx::cols(w, Df, Df), true, M, Df,
                               Epi().gate(Ctx::cols(zc, Df, Df), inv_keep(pd)).pl(dz).colsum(c.grd(p->fam.back().lin.b_off))));
            } else {
                OP(c.gemm(d_cur, false, Ctx::cols(w, 0, Df), true, M, Df,
                               Epi().drop(pd, seed, step, M2F_SITE_MOD(0, 1), Df).pl(dproj[0]).colsum(c.grd(p->mod[0].proj.b_off))));
                OP(c.gemm(d_cur, false, Ctx::cols(w, Df, Df), true, M, Df,
                               Epi().drop(pd, seed, step, M2F_SITE_MOD(1, 1), Df).pl(dproj[1]).colsum(c.grd(p->mod[1].proj.b_off))));
            }
        } else {
            const int mi = p->mod[0].enabled ? 0 : 1;
            OP(c.gemm(d_cur, false, c.pvb(L.w), true, M, Df,
                           Epi().drop(pd, seed, step, M2F_SITE_MOD(mi, 1), Df).pl(dproj[mi]).colsum(c.grd(p->mod[mi].proj.b_off))));
        }
    }
    M2F_TRY(c.done_down_to(md->cls[0].w_off));
    // ---- fusion stack, last layer first
    if (fam) {
        const int F = (int)p->fam.size();
        m2f_planes_t a = Ctx::cols(zc, 0, Df);
        m2f_planes_t dkall = c.pvb(p->dkall_pl), kall = c.pvb(p->kall_pl);
        float* gt = c.f32(p->gt_f32);
        const int64_t gt_ld = p->gt_f32.ld;
        int n = 0;
        for (int f = F - 1; f >= 0; --f, ++n) {
            FusionLayer& fl = p->fam[f];
            // scratch rotates (dz over three buffers, the rest over two); a buffer is rewritten only after the weight
            // gradients of the layer that last read it -- two layers up -- have been computed
            if (n >= 2 && ws1 != c.st) M2F_CUDA(cudaStreamWaitEvent(c.st, p->wgf_ev[n & 1], 0));
            m2f_planes_t dz_cur = c.pvb(p->dz[n % 3]), dz_other = c.pvb(p->dz[(n + 1) % 3]);
            m2f_planes_t dxo = c.pvb(p->dxo[n & 1]), dattn = c.pvb(p->dattn[n & 1]), dqv = c.pvb(p->dqv[n & 1]);
            m2f_planes_t u = c.pvb(fl.u_pl), qv = c.pvb(fl.qv_pl), xo = c.pvb(fl.xo_pl);
            m2f_planes_t t_in = (f == 0) ? c.pvb(p->t0_pl) : Ctx::cols(u, Df, Df);  // the layer's text input
            m2f_planes_t wl = c.pvb(fl.lin.w);
            // linear on the concat
            M2F_TRY(lin_wgrad(c, ws1, dz_cur, u, c.grd(fl.lin.w_off), fl.lin.in, nullptr, fl.lin.out, fl.lin.in));
            // left half of d u -> gradient w.r.t. the out-proj output (gate: relu(x) > 0)
            OP(c.gemm(dz_cur, false, Ctx::cols(wl, 0, Df), true, M, Df, Epi().gate(Ctx::cols(u, 0, Df), 1.f).pl(dxo).colsum(c.grd(fl.wo.b_off))));
            // right half of d u -> first contribution to the gradient of this layer's text input, already in the
            // pre-activation space of its producer: previous fusion linear (relu + dropout: gate [t > 0] / (1-p))
            // or the text projection (cat-relu gate [t0 > 0], then the regenerated projection dropout mask)
            // (needed only by the layer's last GEMM: it runs beside the chain on the audio stack's idle weight-gradient stream)
            const bool gfork = p->use_streams && fusion_fork() && c.side(2) != c.st;
            if (gfork) {
                M2F_TRY(c.dep(c.st, c.side(2)));
                c.cur = c.side(2);
            }
            if (f > 0) {
                OP(c.gemm(dz_cur, false, Ctx::cols(wl, Df, Df), true, M, Df, Epi().gate(t_in, inv_keep(pd)).f32(gt, gt_ld)));
            } else {
                OP(c.gemm(dz_cur, false, Ctx::cols(wl, Df, Df), true, M, Df,
                               Epi().gate(t_in, 1.f).drop(pd, seed, step, M2F_SITE_MOD(1, 1), Df).f32(gt, gt_ld)));
            }
            c.cur = nullptr;
            // attention out-proj
            M2F_TRY(lin_wgrad(c, ws1, dxo, xo, c.grd(fl.wo.w_off), fl.wo.in, nullptr, fl.wo.out, fl.wo.in));
            OP(c.gemm(dxo, false, c.pvb(fl.wo.w), true, M, Df, Epi().pl(dattn)));
            m2f_planes_t q = Ctx::cols(qv, 0, Df), v = Ctx::cols(qv, Df, Df), k = Ctx::cols(kall, (int64_t)f * Df, Df);
            m2f_planes_t dq = Ctx::cols(dqv, 0, Df), dv = Ctx::cols(dqv, Df, Df), dk = Ctx::cols(dkall, (int64_t)f * Df, Df);
            // in-projection weight / bias gradients: rows [0,E) = q, [E,2E) = k, [2E,3E) = v of in_proj_weight
            float* dWin = c.grd(fl.inw_off);
            float* dbin = c.grd(fl.inb_off);
            OP(m2f_attn_bwd(&q, &k, &v, &dattn, key_pad, p->B, p->S, md->cfg.fam_heads, pd, seed, step, M2F_SITE_FAM(f, 0), &dq, &dk, &dv,
                                 dbin, dbin + Df, dbin + 2 * Df, c.stream()));
            M2F_TRY(lin_wgrad(c, ws1, dq, t_in, dWin, Df, nullptr, Df, Df));
            M2F_TRY(lin_wgrad(c, ws1, dv, t_in, dWin + (int64_t)2 * Df * Df, Df, nullptr, Df, Df));
            M2F_TRY(lin_wgrad(c, ws1, dk, a, dWin + (int64_t)Df * Df, Df, nullptr, Df, Df));
            if (ws1 != c.st) M2F_CUDA(cudaEventRecord(p->wgf_ev[n & 1], ws1));
            M2F_TRY(c.done_down_to(fl.inw_off));
            // second + third contributions to the text-input gradient: [dq | dv] . [Wq; Wv], same gate, summed with gt
            if (gfork) M2F_TRY(c.dep(c.side(2), c.st));
            if (f > 0) {
                OP(c.gemm(dqv, false, c.pvb(fl.qv.w), true, M, Df,
                               Epi().gate(t_in, inv_keep(pd)).resid(gt, gt_ld).pl(dz_other).colsum(c.grd(p->fam[f - 1].lin.b_off))));
            } else {
                OP(c.gemm(dqv, false, c.pvb(fl.qv.w), true, M, Df,
                               Epi().drop(pd, seed, step, M2F_SITE_MOD(1, 1), Df).resid(gt, gt_ld).pl(dproj[1]).colsum(c.grd(p->mod[1].proj.b_off))));
            }
        }
        // audio: every layer's K projection + the classifier half, through the projection dropout mask
        OP(c.gemm(dkall, false, c.pvb(md->wk_all.w), true, M, Df,
                       Epi().drop(pd, seed, step, M2F_SITE_MOD(0, 1), Df).resid(c.f32(p->ga_f32), p->ga_f32.ld).pl(dproj[0]).colsum(c.grd(p->mod[0].proj.b_off))));
    }
    // NOTE on the f == 0 text contribution through q/v: q and v read t0 itself (no cat-relu on that path), so only the
    // projection dropout mask applies -- handled above by not gating the [dq | dv] GEMM at f == 0.
    // text first, audio last: gradients then complete in exactly the reverse of the arena order
    // The two stacks are independent from here on: text continues on the caller's stream, audio runs beside it.
    c.narrow = narrow_level() >= 2;  // the two stacks run side by side from here on
    const bool interleave = (c.cb || c.marks || c.xchg) && p->mod[0].enabled && p->mod[1].enabled;
    if (interleave) {
        M2F_TRY(c.done_down_to(fam ? p->fam[0].inw_off : md->cls[0].w_off, true));  // everything above the stacks
        c.collect_begin(1, ws1);
    }
    // fork before the text stack is enqueued: the audio chain needs only what the main stream holds so far (its dproj)
    cudaStream_t as = p->mod[1].enabled ? c.side(0) : c.st;
    if (p->mod[0].enabled) M2F_TRY(c.dep(c.st, as));
    if (p->mod[1].enabled) M2F_TRY(backward_modality(c, p->mod[1], key_pad, dproj[1], ws1));
    if (interleave) {
        M2F_TRY(c.done_down_to(p->mod[1].layers[0].qkv.w_off, true));
        c.collect_begin(0, c.side(2));
    }
    if (p->mod[0].enabled) {
        c.cur = (as == c.st) ? nullptr : as;
        int rc = backward_modality(c, p->mod[0], key_pad, dproj[0], c.side(2));
        c.cur = nullptr;
        M2F_TRY(rc);
    }
    if (interleave) {
        M2F_TRY(c.done_down_to(0, true));
        c.collect_end();
        M2F_TRY(c.report_stacks());
    }
    M2F_TRY(c.done_down_to(0, true));
    for (int i = 0; i < 3; ++i) M2F_TRY(c.dep(c.side(i), c.st));
    if (c.cb || c.marks || c.xchg) {  // biases / LayerNorm parameters (atomically accumulated by every stream): final once everything has joined
        cudaStream_t ready = p->use_streams ? p->ready : c.st;
        M2F_TRY(c.dep(c.st, ready));
        M2F_TRY(c.report(md->off_1d, md->arena_floats, ready, true));
        if (c.tail_adam) M2F_TRY(c.adam_tail());  // range by range behind the exchange, on the caller's stream
        // a capture must end with every forked stream joined; with the exchange inside the backward the caller's stream
        // continues (optimizer) only after the last range has been exchanged
        if ((c.marks || c.xchg) && c.ready_in_capture) M2F_TRY(c.dep(ready, c.st));
    }
    return M2F_OK;
}

}  // namespace m2f

// ------------------------------------------------------------------------------------------ C ABI
extern "C" int m2f_model_create(const m2f_model_cfg_t* cfg, m2f_model_t** out) {
    M2F_REQUIRE(cfg && out, M2F_ERR_ARG, "model_create: null argument");
    m2f_model* md = new m2f_model();
    md->cfg = *cfg;
    const int rc = build_model(md);
    if (rc != M2F_OK) {
        delete md;
        return rc;
    }
    *out = md;
    return M2F_OK;
}
extern "C" int m2f_model_destroy(m2f_model_t* m) {
    delete m;
    return M2F_OK;
}
extern "C" int64_t m2f_model_num_params(const m2f_model_t* m) { return m ? (int64_t)m->params.size() : 0; }
extern "C" int m2f_model_param_info(const m2f_model_t* m, int64_t i, char* name_buf, int32_t name_buf_len, int64_t* offset, int64_t* rows,
                                    int64_t* cols) {
    M2F_REQUIRE(m && i >= 0 && i < (int64_t)m->params.size(), M2F_ERR_ARG, "param_info: index out of range");
    const ParamInfo& pi = m->params[(size_t)i];
    if (name_buf && name_buf_len > 0) snprintf(name_buf, (size_t)name_buf_len, "%s", pi.name.c_str());
    if (offset) *offset = pi.off;
    if (rows) *rows = pi.rows;
    if (cols) *cols = pi.cols;
    return M2F_OK;
}
extern "C" int64_t m2f_model_arena_floats(const m2f_model_t* m) { return m ? m->arena_floats : 0; }
extern "C" int64_t m2f_model_arena_1d_offset(const m2f_model_t* m) { return m ? m->off_1d : 0; }
extern "C" int64_t m2f_model_shadow_bytes(const m2f_model_t* m) { return m ? m->shadow_bytes : 0; }

extern "C" int m2f_model_refresh_shadow(const m2f_model_t* m, const float* params, void* shadow, void* stream) {
    M2F_REQUIRE(m && params && shadow, M2F_ERR_ARG, "refresh_shadow: null argument");
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(params) % 16) == 0 && (reinterpret_cast<uintptr_t>(shadow) % 256) == 0, M2F_ERR_ALIGN,
                "refresh_shadow: params must be 16-byte and shadow 256-byte aligned");
    auto st = reinterpret_cast<cudaStream_t>(stream);
    for (size_t i = 0; i < m->sentries.size(); i += SHADOW_CHUNK) {
        ShadowBatch b;
        b.n = (int)std::min<size_t>(SHADOW_CHUNK, m->sentries.size() - i);
        for (int j = 0; j < b.n; ++j) b.e[j] = m->sentries[i + j];
        shadow_refresh_kernel<<<dim3(64, b.n), 256, 0, st>>>(params, reinterpret_cast<uint8_t*>(shadow), b);
        M2F_LAUNCH_CHECK();
    }
    return M2F_OK;
}

extern "C" int m2f_model_adam_step(const m2f_model_t* m, float* params, const float* grads, float* exp_avg, float* exp_avg_sq, void* shadow,
                                   float lr, float beta1, float beta2, float eps, float weight_decay, int64_t t, float grad_scale,
                                   void* stream) {
    M2F_REQUIRE(m && params && grads && exp_avg && exp_avg_sq && shadow && t >= 1, M2F_ERR_ARG, "model_adam_step: bad argument");
    M2F_REQUIRE(((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) | reinterpret_cast<uintptr_t>(exp_avg) |
                  reinterpret_cast<uintptr_t>(exp_avg_sq)) % 16) == 0 && (reinterpret_cast<uintptr_t>(shadow) % 256) == 0,
                M2F_ERR_ALIGN, "model_adam_step: arenas must be 16-byte, shadow 256-byte aligned");
    auto st = reinterpret_cast<cudaStream_t>(stream);
    const double bc1 = 1.0 - pow((double)beta1, (double)t), bc2 = 1.0 - pow((double)beta2, (double)t);
    AdamArgs a{beta1, beta2, eps, weight_decay, (float)((double)lr / bc1), (float)(1.0 / sqrt(bc2)), grad_scale};
    // 2-D weights: every element belongs to exactly one planes entry of the shadow table
    std::vector<SEntry> copies;
    ShadowBatch b;
    b.n = 0;
    auto flush = [&]() -> int {
        if (b.n == 0) return M2F_OK;
        adam_shadow_kernel<<<dim3(96, b.n), 256, 0, st>>>(params, grads, exp_avg, exp_avg_sq, reinterpret_cast<uint8_t*>(shadow), b, a, nullptr);
        M2F_LAUNCH_CHECK();
        b.n = 0;
        return M2F_OK;
    };
    for (const SEntry& e : m->sentries) {
        if (e.planes == 0) {
            copies.push_back(e);
            continue;
        }
        b.e[b.n++] = e;
        if (b.n == SHADOW_CHUNK) M2F_TRY(flush());
    }
    M2F_TRY(flush());
    // biases / LayerNorm vectors, then their packed fp32 copies in the shadow
    M2F_TRY(m2f_adam_step(params + m->off_1d, grads + m->off_1d, exp_avg + m->off_1d, exp_avg_sq + m->off_1d, m->arena_floats - m->off_1d, lr,
                          beta1, beta2, eps, weight_decay, t, grad_scale, stream));
    for (size_t i = 0; i < copies.size(); i += SHADOW_CHUNK) {
        b.n = (int)std::min<size_t>(SHADOW_CHUNK, copies.size() - i);
        for (int j = 0; j < b.n; ++j) b.e[j] = copies[i + j];
        shadow_refresh_kernel<<<dim3(4, b.n), 256, 0, st>>>(params, reinterpret_cast<uint8_t*>(shadow), b);
        M2F_LAUNCH_CHECK();
    }
    return M2F_OK;
}

extern "C" int m2f_plan_create(m2f_model_t* m, int64_t B, int64_t S, int32_t training_plan, m2f_plan_t** out) {
    M2F_REQUIRE(m && out && B > 0 && S > 0, M2F_ERR_ARG, "plan_create: bad argument");
    M2F_REQUIRE(B * S < (1ll << 30), M2F_ERR_UNSUPPORTED, "plan_create: B*S too large");
    m2f_plan* p = new m2f_plan();
    p->m = m;
    p->B = B; p->S = S; p->M = B * S;
    p->training = training_plan;
    const int64_t M = p->M;
    const int P = m->cfg.planes, FF = m->cfg.ffn_dim, Df = m->Df;
    const int Pb = m->cfg.planes_bwd > 0 ? m->cfg.planes_bwd : P;  // planes of the backward's own buffers
    int Dmax = 0;
    p->mod[0] = m->mod[0];
    p->mod[1] = m->mod[1];
    p->fam = m->fam;
    {
        Buf kb;
        kb.rows = 1; kb.cols = M; kb.ld = M; kb.planes = 0;
        kb.off = p->ws_bytes;
        p->ws_bytes += round_up(M, 256);
        p->key_pad = kb;
    }
    // Inference plans recycle buffers (training plans keep every activation for the backward):
    //  * the encoder layers of a stack share one set of per-layer buffers, and inside that set tensors with disjoint lifetimes
    //    share storage (every kernel of a stack runs in stream order): h (FFN hidden) lies on qkv, x1_pl on o_pl, y2 on y1;
    //  * layer outputs alternate between two slots; the fp32 copy of the last stack output is never written and not allocated;
    //  * the fusion layers share one q|v, one attention-output and two alternating concat buffers, and these, the stacked K
    //    projections and the classifier's hidden layers are laid over the audio stack's buffers: everything written after the
    //    two stacks have joined, when no stack buffer is live any more (their last reader is the stack's projection GEMM).
    // C5 (M = 131 072, default dims): 22 GB -> 7.4 GB in bf16 mode, 48 GB -> 16.6 GB with three planes.
    auto overlay = [&](auto first, auto second) {  // second() is laid over first(): both start at the same offset
        p->ws_bytes = round_up(p->ws_bytes, 256);
        const int64_t mark = p->ws_bytes;
        first();
        const int64_t end_first = p->ws_bytes;
        p->ws_bytes = mark;
        second();
        p->ws_bytes = std::max(p->ws_bytes, end_first);
    };
    int64_t stack_lo[2] = {0, 0}, stack_hi[2] = {0, 0};
    for (int i = 0; i < 2; ++i) {
        Modality& mo = p->mod[i];
        if (!mo.enabled) continue;
        const int D = mo.D;
        Dmax = std::max(Dmax, D);
        p->ws_bytes = round_up(p->ws_bytes, 256);
        stack_lo[i] = p->ws_bytes;
        mo.in_pl = p->alloc_pl(M, D, P);
        EncLayer shared;  // inference plans alias one set of per-layer buffers
        bool have_shared = false;
        size_t li = 0;
        Buf x2_slot[2], x2pl_slot[2];
        for (auto& e : mo.layers) {
            if (training_plan) {
                e.qkv_pl = p->alloc_pl(M, 3 * D, P);
                e.o_pl = p->alloc_pl(M, D, P);
                e.y1 = p->alloc_f32(M, D);
                e.st1 = p->alloc_f32(M, 2);
                e.x1 = p->alloc_f32(M, D);
                e.x1_pl = p->alloc_pl(M, D, P);
                e.h_pl = p->alloc_pl(M, FF, P);
                e.y2 = p->alloc_f32(M, D);
                e.st2 = p->alloc_f32(M, 2);
            } else if (!have_shared) {
                overlay([&] { e.qkv_pl = p->alloc_pl(M, 3 * D, P); }, [&] { e.h_pl = p->alloc_pl(M, FF, P); });
                overlay([&] { e.o_pl = p->alloc_pl(M, D, P); }, [&] { e.x1_pl = p->alloc_pl(M, D, P); });
                overlay([&] { e.y1 = p->alloc_f32(M, D); }, [&] { e.y2 = p->alloc_f32(M, D); });
                e.st1 = p->alloc_f32(M, 2);
                e.x1 = p->alloc_f32(M, D);
                e.st2 = p->alloc_f32(M, 2);
                shared = e;
                have_shared = true;
            } else {
                e.qkv_pl = shared.qkv_pl; e.o_pl = shared.o_pl; e.y1 = shared.y1; e.st1 = shared.st1; e.x1 = shared.x1;
                e.x1_pl = shared.x1_pl; e.h_pl = shared.h_pl; e.y2 = shared.y2; e.st2 = shared.st2;
            }
            // the layer output feeds the next layer while this layer's buffers are being reused: never aliased per layer pair
            if (training_plan || li < 2) {
                e.x2 = p->alloc_f32(M, D);
                e.x2_pl = p->alloc_pl(M, D, P);
                if (li < 2) { x2_slot[li] = e.x2; x2pl_slot[li] = e.x2_pl; }
            } else {  // two alternating output slots are enough for inference
                e.x2 = x2_slot[li & 1];
                e.x2_pl = x2pl_slot[li & 1];
            }
            ++li;
        }
        mo.xout.resize(mo.T); mo.xout_pl.resize(mo.T); mo.stf.resize(mo.T);
        for (int t = 0; t < mo.T; ++t) {
            if (training_plan || t < mo.T - 1) mo.xout[t] = p->alloc_f32(M, D);
            else { mo.xout[t] = Buf(); mo.xout[t].rows = M; mo.xout[t].cols = D; mo.xout[t].ld = round_up(D, 4); mo.xout[t].off = stack_lo[i]; }  // never written (forward_modality)
            mo.xout_pl[t] = p->alloc_pl(M, D, P);
            mo.stf[t] = p->alloc_f32(M, 2);
        }
        p->ws_bytes = round_up(p->ws_bytes, 256);
        stack_hi[i] = p->ws_bytes;
    }
    p->zc_pl = p->alloc_pl(M, m->cls_in, P);
    if (!m->fam.empty()) p->t0_pl = p->alloc_pl(M, Df, P);
    p->cls_h.resize(m->cls.size() - 1);
    auto after_join = [&] {  // buffers first written after the two stacks have joined
        if (!m->fam.empty()) {
            const int F = (int)m->fam.size();
            p->kall_pl = p->alloc_pl(M, (int64_t)F * Df, P);
            int fi = 0;
            for (auto& fl : p->fam) {
                if (training_plan || fi == 0) {
                    fl.qv_pl = p->alloc_pl(M, 2 * Df, P);
                    fl.xo_pl = p->alloc_pl(M, Df, P);
                } else {
                    fl.qv_pl = p->fam[0].qv_pl;
                    fl.xo_pl = p->fam[0].xo_pl;
                }
                // layer f reads its concat operand u[f] and writes the text half of u[f + 1]
                fl.u_pl = (training_plan || fi < 2) ? p->alloc_pl(M, 2 * Df, P) : p->fam[fi & 1].u_pl;
                ++fi;
            }
        }
        for (size_t j = 0; j + 1 < m->cls.size(); ++j) p->cls_h[j] = p->alloc_pl(M, m->cls[j].out, P);
    };
    if (!training_plan && !m->fam.empty() && p->mod[0].enabled && stack_hi[0] > stack_lo[0]) {
        const int64_t save = p->ws_bytes;
        p->ws_bytes = stack_lo[0];
        after_join();
        if (p->ws_bytes > stack_hi[1] && p->ws_bytes > stack_hi[0]) {  // larger than the stacks' region (tiny encoders): no overlay
            p->ws_bytes = save;
            after_join();
        } else {
            p->ws_bytes = save;
        }
    } else {
        after_join();
    }
    if (training_plan) {
        p->dl_pl = p->alloc_pl(M, 8 * ((m->cfg.cls_out + 7) / 8), Pb);
        // every backward scratch matrix gets ld = its own width; the two stacks run side by side, so each has its own
        for (int i = 0; i < 2; ++i) {
            if (!m->mod[i].enabled) continue;
            const int D = m->mod[i].D;
            m2f_plan::StackScratch& sc = p->ss[i];
            sc.gG = p->alloc_f32(M, D);
            sc.gA = p->alloc_f32(M, D);
            sc.gB = p->alloc_f32(M, D);
            for (int j = 0; j < 2; ++j) {
                sc.dx2[j] = p->alloc_pl(M, D, Pb);
                sc.dx1[j] = p->alloc_pl(M, D, Pb);
                sc.dh[j] = p->alloc_pl(M, FF, Pb);
                sc.do_[j] = p->alloc_pl(M, D, Pb);
                sc.dqkv[j] = p->alloc_pl(M, 3 * D, Pb);
            }
        }
        p->dcls_a = p->alloc_pl(M, m->cfg.cls_hidden, Pb);
        p->dcls_b = p->alloc_pl(M, m->cfg.cls_hidden, Pb);
        for (int j = 0; j < 3; ++j) p->dz[j] = p->alloc_pl(M, Df, Pb);
        p->dproj_pl = p->alloc_pl(M, Df, Pb);
        p->dprojt_pl = p->alloc_pl(M, Df, Pb);
        if (!m->fam.empty()) {
            for (int j = 0; j < 2; ++j) {
                p->dxo[j] = p->alloc_pl(M, Df, Pb);
                p->dattn[j] = p->alloc_pl(M, Df, Pb);
                p->dqv[j] = p->alloc_pl(M, 2 * Df, Pb);
            }
            p->dkall_pl = p->alloc_pl(M, (int64_t)m->fam.size() * Df, Pb);
            p->gt_f32 = p->alloc_f32(M, Df);
            p->ga_f32 = p->alloc_f32(M, Df);
        }
    }
    if (const char* e = std::getenv("M2F_STREAMS")) p->use_streams = (e[0] != '0');
    if (const char* e = std::getenv("M2F_GRAPH")) p->use_graph = (e[0] != '0');      // 0: launch every kernel eagerly
    if (m->mod[1].enabled) p->text_in = p->alloc_f32(M, m->mod[1].D);
    if (m->mod[0].enabled) p->audio_in = p->alloc_f32(M, m->mod[0].D);  // 0: everything on the caller's stream
    p->ws_bytes = round_up(p->ws_bytes, 256);
    *out = p;
    return M2F_OK;
}
extern "C" int m2f_plan_destroy(m2f_plan_t* p) {
    if (p) p->release_streams();
    delete p;
    return M2F_OK;
}
extern "C" int64_t m2f_plan_workspace_bytes(const m2f_plan_t* p) { return p ? p->ws_bytes : 0; }

static int bind(m2f_plan* p, const void* shadow, void* ws) {
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(ws) % 256) == 0 && (reinterpret_cast<uintptr_t>(shadow) % 256) == 0, M2F_ERR_ALIGN,
                "plan: workspace and shadow must be 256-byte aligned");
    if (p->bound_ws != ws || p->bound_shadow != shadow) {
        p->tmaps.clear();
        p->epimaps.clear();
        p->bound_ws = ws;
        p->bound_shadow = shadow;
    }
    return M2F_OK;
}

// Capture `body` (enqueued on the plan's capture stream, forking onto its side streams) once per key, then replay it
// into the caller's stream.  Dropout kernels captured here read (seed, step) from p->rng_dev.
template <class F>
static int run_graphed(m2f_plan* p, m2f_plan::GraphSlot& g, const m2f_plan::GraphKey& key, cudaStream_t user, F&& body) {
    if (!g.exec || !(g.key == key)) {
        g.reset();
        const long long l0 = m2f_launch_count();
        M2F_CUDA(cudaStreamBeginCapture(p->cap, cudaStreamCaptureModeThreadLocal));
        m2f_set_rng_source(reinterpret_cast<const uint64_t*>(p->rng_dev));
        const int rc = body(p->cap);
        m2f_set_rng_source(nullptr);
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(p->cap, &graph);
        const long long captured = m2f_launch_count() - l0;
        note_launches(-captured);  // nothing ran yet
        if (rc != M2F_OK) {
            if (graph) cudaGraphDestroy(graph);
            return rc;
        }
        M2F_CUDA(e);
        const cudaError_t ei = cudaGraphInstantiate(&g.exec, graph, 0);
        cudaGraphDestroy(graph);
        M2F_CUDA(ei);
        g.key = key;
        g.kernels = captured;
    }
    static const bool host_time = std::getenv("M2F_HOSTTIME") != nullptr;
    if (host_time) {
        const auto t0 = std::chrono::steady_clock::now();
        M2F_CUDA(cudaGraphLaunch(g.exec, user));
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
        std::fprintf(stderr, "m2f: cudaGraphLaunch of %lld kernels took %.1f us on the host\n", g.kernels, us);
    } else {
        M2F_CUDA(cudaGraphLaunch(g.exec, user));
    }
    note_launches(g.kernels);
    return M2F_OK;
}

__global__ void set_rng_kernel(unsigned long long* dst, unsigned long long seed, unsigned long long step) {
    dst[0] = seed;
    dst[1] = step;
}

extern "C" int m2f_plan_forward(m2f_plan_t* p, const float* params, const void* shadow, void* workspace, const float* text,
                                const float* audio, const uint8_t* key_pad, int32_t dropout_on, uint64_t seed, uint64_t step, float* logits,
                                void* stream) {
    M2F_REQUIRE(p && params && shadow && workspace && key_pad && logits, M2F_ERR_ARG, "plan_forward: null argument");
    m2f_model* md = p->m;
    M2F_REQUIRE(!md->mod[0].enabled || audio, M2F_ERR_ARG, "plan_forward: audio is enabled but the audio pointer is NULL");
    M2F_REQUIRE(!md->mod[1].enabled || text, M2F_ERR_ARG, "plan_forward: text is enabled but the text pointer is NULL");
    M2F_REQUIRE((!audio || (reinterpret_cast<uintptr_t>(audio) % 16) == 0) && (!text || (reinterpret_cast<uintptr_t>(text) % 16) == 0) &&
                    (reinterpret_cast<uintptr_t>(logits) % 16) == 0 && (reinterpret_cast<uintptr_t>(params) % 16) == 0,
                M2F_ERR_ALIGN, "plan_forward: text / audio / logits / params must be 16-byte aligned");
    M2F_TRY(bind(p, shadow, workspace));
    M2F_TRY(p->ensure_streams());
    p->p_used = dropout_on ? md->cfg.dropout : 0.f;
    p->seed = seed;
    p->step = step;
    p->have_loss_grad = false;
    cudaStream_t user = reinterpret_cast<cudaStream_t>(stream);
    Ctx c{p, params, nullptr, reinterpret_cast<uint8_t*>(const_cast<void*>(shadow)), reinterpret_cast<uint8_t*>(workspace), user,
          md->cfg.planes};
    uint8_t* kp = c.ws + p->key_pad.off;
    M2F_CUDA(cudaMemcpyAsync(kp, key_pad, (size_t)p->M, cudaMemcpyDeviceToDevice, user));
    const bool timeline = std::getenv("M2F_TIMELINE") != nullptr && p->training;
    if (timeline) {
        if (!p->tl_dev) M2F_CUDA(cudaMalloc(&p->tl_dev, TL_CAP * sizeof(unsigned long long)));
        p->tl_on = true;
    }
    if (!p->use_graph) {
        if (timeline) p->tl_labels.clear();
        return forward_all(c, text, audio, kp, logits, PH_ALL);
    }
    // graph mode: stage what changes per call (inputs, rng counters) into plan-owned memory, replay the body, finish eagerly
    const float* t_in = nullptr;
    const float* a_in = nullptr;
    if (md->mod[1].enabled) {
        M2F_CUDA(cudaMemcpyAsync(c.f32(p->text_in), text, (size_t)p->M * md->mod[1].D * 4, cudaMemcpyDeviceToDevice, user));
        t_in = c.f32(p->text_in);
    }
    if (md->mod[0].enabled) {
        M2F_CUDA(cudaMemcpyAsync(c.f32(p->audio_in), audio, (size_t)p->M * md->mod[0].D * 4, cudaMemcpyDeviceToDevice, user));
        a_in = c.f32(p->audio_in);
    }
    set_rng_kernel<<<1, 1, 0, user>>>(p->rng_dev, seed, step);
    M2F_LAUNCH_CHECK();
    m2f_plan::GraphKey key;
    key.params = params; key.shadow = shadow; key.ws = workspace; key.p_used = p->p_used;
    if (timeline && (!p->gf.exec || !(p->gf.key == key))) {  // stamps are captured with the graph: slots restart at 0
        p->tl_labels.clear();
        p->gb.reset();
    }
    M2F_TRY(run_graphed(p, p->gf, key, user, [&](cudaStream_t cs) {
        Ctx g = c;
        g.st = cs;
        return forward_all(g, t_in, a_in, kp, nullptr, PH_BODY);
    }));
    return forward_all(c, t_in, a_in, kp, logits, PH_TAIL);
}

extern "C" int m2f_plan_loss(m2f_plan_t* p, void* workspace, const float* logits, const int64_t* target, const float* class_w, float eps,
                             float grad_scale, float* loss_out, void* stream) {
    M2F_REQUIRE(p && workspace && logits && target && loss_out, M2F_ERR_ARG, "plan_loss: null argument");
    M2F_REQUIRE(p->training, M2F_ERR_ARG, "plan_loss: needs a training plan");
    Ctx c{p, nullptr, nullptr, nullptr, reinterpret_cast<uint8_t*>(workspace), reinterpret_cast<cudaStream_t>(stream), p->m->cfg.planes};
    m2f_planes_t dl = c.pv(p->dl_pl);
    M2F_TRY(m2f_ce_fwd_bwd(logits, p->m->cfg.cls_out, target, class_w, p->M, p->m->cfg.cls_out, eps, grad_scale, loss_out, nullptr, 0, &dl,
                           stream));
    p->have_loss_grad = true;
    return M2F_OK;
}

extern "C" int m2f_plan_set_exchange(m2f_plan_t* p, const m2f_peer_group_t* group, float scale, int32_t ctas) {
    M2F_REQUIRE(p, M2F_ERR_ARG, "plan_set_exchange: plan is NULL");
    p->gb.reset();  // the captured backward embeds (or lacks) the exchange kernels
    if (!group) {
        p->xchg_on = false;
        return M2F_OK;
    }
    M2F_REQUIRE(group->world >= 1 && group->world <= M2F_XCHG_MAX_RANKS && group->rank >= 0 && group->rank < group->world, M2F_ERR_ARG,
                "plan_set_exchange: rank %d / world %d outside [1, %d]", group->rank, group->world, M2F_XCHG_MAX_RANKS);
    p->xchg_group = *group;
    p->xchg_scale = scale;
    p->xchg_ctas = ctas;
    p->xchg_on = true;
    return M2F_OK;
}

extern "C" int64_t m2f_plan_exchange_ranges(const m2f_plan_t* p) { return p ? p->xchg_ranges : 0; }

extern "C" int m2f_plan_set_optimizer(m2f_plan_t* p, float* exp_avg, float* exp_avg_sq, const float* hyper_dev) {
    M2F_REQUIRE(p, M2F_ERR_ARG, "plan_set_optimizer: plan is NULL");
    p->gb.reset();  // the captured backward embeds (or lacks) the update kernels
    if (!exp_avg) {
        p->opt_on = false;
        return M2F_OK;
    }
    M2F_REQUIRE(exp_avg_sq && hyper_dev, M2F_ERR_ARG, "plan_set_optimizer: null argument");
    M2F_REQUIRE(((reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq) | reinterpret_cast<uintptr_t>(hyper_dev)) % 16) == 0,
                M2F_ERR_ALIGN, "plan_set_optimizer: state and hyper-parameter buffers must be 16-byte aligned");
    M2F_REQUIRE(p->training, M2F_ERR_ARG, "plan_set_optimizer: needs a training plan");
    p->opt_m = exp_avg;
    p->opt_v = exp_avg_sq;
    p->opt_hp = hyper_dev;
    p->opt_on = true;
    return M2F_OK;
}

extern "C" int m2f_adam_hyper(float* hyper_dev, float lr, float beta1, float beta2, float eps, float weight_decay, int64_t t,
                              float grad_scale, void* stream) {
    M2F_REQUIRE(hyper_dev && t >= 1, M2F_ERR_ARG, "adam_hyper: bad argument");
    static_assert(sizeof(AdamArgs) <= M2F_ADAM_HYPER_FLOATS * sizeof(float), "hyper-parameter block too small");
    const double bc1 = 1.0 - pow((double)beta1, (double)t), bc2 = 1.0 - pow((double)beta2, (double)t);
    const AdamArgs a{beta1, beta2, eps, weight_decay, (float)((double)lr / bc1), (float)(1.0 / sqrt(bc2)), grad_scale};
    adam_hyper_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(reinterpret_cast<AdamArgs*>(hyper_dev), a);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_plan_backward_bucketed(m2f_plan_t* p, const float* params, const void* shadow, void* workspace,
                                          const float* dlogits, float* grads, int64_t bucket_floats, m2f_bucket_cb cb, void* user,
                                          void* stream) {
    M2F_REQUIRE(p && params && shadow && workspace && grads, M2F_ERR_ARG, "plan_backward: null argument");
    M2F_REQUIRE(p->training, M2F_ERR_ARG, "plan_backward: this plan was created for inference (training_plan = 0)");
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(grads) % 16) == 0, M2F_ERR_ALIGN, "plan_backward: grads must be 16-byte aligned");
    M2F_TRY(bind(p, shadow, workspace));
    M2F_TRY(p->ensure_streams());
    cudaStream_t ust = reinterpret_cast<cudaStream_t>(stream);
    Ctx c{p, params, grads, reinterpret_cast<uint8_t*>(const_cast<void*>(shadow)), reinterpret_cast<uint8_t*>(workspace), ust,
          p->m->cfg.planes};
    c.bucket_floats = bucket_floats > 0 ? bucket_floats : 1;
    c.cb = cb;
    c.cb_user = user;
    c.hi = p->m->off_1d;
    c.xchg = !cb && bucket_floats > 0 && p->xchg_on;
    c.tail_adam = c.xchg && p->opt_on;
    if (c.xchg) {
        M2F_REQUIRE(grads == p->xchg_group.peer_base[p->xchg_group.rank], M2F_ERR_ARG,
                    "plan_backward: with an exchange attached the gradients must be written to this rank's copy of the symmetric buffer");
        M2F_REQUIRE(p->xchg_group.signal_off_bytes >= p->m->arena_floats * 4, M2F_ERR_ARG, "plan_backward: the exchange's signal area overlaps the gradient arena");
        p->xchg_ranges = 0;
    }
    const uint8_t* kp = c.ws + p->key_pad.off;
    int rc;
    if (!p->use_graph) {
        rc = backward_all(c, kp, dlogits, PH_ALL);
    } else {
        // With bucket callbacks the graph carries one external event-record node per reported range; after the launch each
        // callback gets the plan's ready stream, ordered behind its event, so the caller's collectives still overlap the rest.
        M2F_TRY(backward_all(c, kp, dlogits, PH_HEAD));
        m2f_plan::GraphKey key;
        key.params = params; key.shadow = shadow; key.ws = workspace; key.grads = grads; key.p_used = p->p_used;
        key.flags = (cb ? 1 : (c.xchg ? 2 : 0)) | (c.tail_adam ? 4 : 0);
        key.opt_m = c.tail_adam ? p->opt_m : nullptr;
        key.bucket_floats = (cb || c.xchg) ? c.bucket_floats : 0;
        std::vector<std::pair<int64_t, int64_t>> marks;
        bool captured = false;
        rc = run_graphed(p, p->gb, key, ust, [&](cudaStream_t cs) {
            Ctx g = c;
            g.st = cs;
            g.cb = nullptr;
            g.capturing = true;
            if (cb) g.marks = &marks;
            if (c.xchg) p->xchg_ranges = 0;
            captured = true;
            int r = backward_all(g, kp, nullptr, PH_BODY);
            if (r == M2F_OK) g.stamp("end");
            return r;
        });
        if (rc == M2F_OK && c.xchg) {
            if (captured) p->gb.xchg_ranges = p->xchg_ranges;
            else p->xchg_ranges = p->gb.xchg_ranges;
        }
        if (rc == M2F_OK && cb) {
            if (captured) p->gb.marks = marks;
            for (size_t k = 0; k < p->gb.marks.size(); ++k) {
                M2F_CUDA(cudaStreamWaitEvent(p->ready, p->bucket_evs[k], 0));
                cb(user, p->gb.marks[k].first, p->gb.marks[k].second, p->ready);
            }
        }
    }
    if (p->tl_on && rc == M2F_OK) {
        if (!p->use_graph) c.stamp("end");
        const char* n = std::getenv("M2F_TIMELINE_STEP");
        if (++p->tl_calls == (n ? std::atoll(n) : 5)) {
            M2F_CUDA(cudaDeviceSynchronize());
            std::vector<unsigned long long> t(p->tl_labels.size());
            M2F_CUDA(cudaMemcpy(t.data(), p->tl_dev, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
            if (FILE* f = std::fopen(std::getenv("M2F_TIMELINE"), "w")) {
                std::fprintf(f, "#t0_ns\t%llu\n", t.empty() ? 0ull : t[0]);  // absolute %globaltimer of entry 0
                for (size_t i = 0; i < t.size(); ++i)
                    std::fprintf(f, "%zu\t%d\t%.3f\t%s\n", i, p->tl_labels[i].second, ((double)t[i] - (double)t[0]) * 1e-3,
                                 p->tl_labels[i].first.c_str());
                std::fclose(f);
            }
        }
    }
    return rc;
}

extern "C" int m2f_plan_backward(m2f_plan_t* p, const float* params, const void* shadow, void* workspace, const float* dlogits,
                                 float* grads, void* stream) {
    return m2f_plan_backward_bucketed(p, params, shadow, workspace, dlogits, grads, 0, nullptr, nullptr, stream);
}
