// Either side of the model step:
//
//  * m2f_collate_dialogues - what Dataset.__getitem__ + collate_fn + apply_padding produce on the host
//    (src/dataset.py:32-65,71-89; src/utils.py:15-31), built on the device from embedding tables that stay resident in
//    HBM: gather the utterance rows of B dialogues into [B, S, D] padded with 0, labels padded with -1, mask = (label
//    == -1).  Pure HBM traffic: (Dt + Da) * 4 bytes read and written per utterance, written only for padding.
//  * m2f_batch_metrics - the per-batch accuracy_score / f1_score(average='weighted') of src/test.py:65-71 and
//    src/train.py:262-267 from an on-device confusion matrix, accumulated into running sums so that the host reads
//    three doubles once per epoch instead of predictions and labels once per batch.
#include "common.cuh"

namespace m2f {

constexpr int COLLATE_THREADS = 128;
constexpr int MET_THREADS = 256;
constexpr int MET_MAX_C = 32;

// One CTA per (dialogue slot b, position s).  VEC: rows are 16-byte aligned and widths are multiples of 4.
template <bool VEC>
__global__ void __launch_bounds__(COLLATE_THREADS)
collate_kernel(const float* __restrict__ text_tab, int64_t text_ld, int Dt, const float* __restrict__ audio_tab, int64_t audio_ld, int Da,
               const int64_t* __restrict__ labels, const int64_t* __restrict__ rows, const int64_t* __restrict__ offsets,
               const int64_t* __restrict__ batch, int S, float* __restrict__ text_out, float* __restrict__ audio_out,
               int64_t* __restrict__ emotion_out, uint8_t* __restrict__ mask_out) {
    const int64_t slot = blockIdx.x;
    const int64_t b = slot / S;
    const int s = (int)(slot - b * S);
    const int64_t d = batch[b];
    const int64_t lo = offsets[d], len = offsets[d + 1] - lo;
    const bool live = s < len;
    const int64_t r = live ? rows[lo + s] : 0;
    if (threadIdx.x == 0) {
        emotion_out[slot] = live ? labels[r] : -1;
        mask_out[slot] = live ? 0 : 1;
    }
    float* to = text_out + slot * Dt;
    float* ao = audio_out + slot * Da;
    if (VEC) {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* ts = reinterpret_cast<const float4*>(text_tab + r * text_ld);
        const float4* as = reinterpret_cast<const float4*>(audio_tab + r * audio_ld);
        for (int i = threadIdx.x; i < Dt / 4; i += COLLATE_THREADS) reinterpret_cast<float4*>(to)[i] = live ? __ldg(ts + i) : z;
        for (int i = threadIdx.x; i < Da / 4; i += COLLATE_THREADS) reinterpret_cast<float4*>(ao)[i] = live ? __ldg(as + i) : z;
    } else {
        for (int i = threadIdx.x; i < Dt; i += COLLATE_THREADS) to[i] = live ? text_tab[r * text_ld + i] : 0.f;
        for (int i = threadIdx.x; i < Da; i += COLLATE_THREADS) ao[i] = live ? audio_tab[r * audio_ld + i] : 0.f;
    }
}

// confusion[y * C + argmax(logits row)] += 1 over rows with y != ignore_index.  argmax takes the first maximum
// (torch.argmax on ties, src/test.py:65); a NaN logit wins, as in torch.
__global__ void __launch_bounds__(MET_THREADS)
confusion_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target, int64_t M, int C, int64_t ignore_index,
                 int* __restrict__ confusion, int64_t* __restrict__ pred) {
    __shared__ int cm[MET_MAX_C * MET_MAX_C];
    for (int i = threadIdx.x; i < C * C; i += MET_THREADS) cm[i] = 0;
    __syncthreads();
    for (int64_t m = blockIdx.x * (int64_t)MET_THREADS + threadIdx.x; m < M; m += (int64_t)gridDim.x * MET_THREADS) {
        const float* z = logits + m * ld;
        int best = 0;
        float bv = z[0];
        for (int c = 1; c < C; ++c) {
            const float v = z[c];
            if (!(bv != bv) && (v > bv || v != v)) { bv = v; best = c; }
        }
        if (pred) pred[m] = best;
        const int64_t y = target[m];
        if (y == ignore_index || y < 0 || y >= C) continue;
        atomicAdd(&cm[(int)y * C + best], 1);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C * C; i += MET_THREADS)
        if (cm[i]) atomicAdd(confusion + i, cm[i]);
}

// One warp: lane c owns class c.  sklearn semantics (restated in oracle/m2fnet_oracle.py:batch_metrics): classes
// present in y_true or y_pred, F1_c = 2 tp / (2 tp + fp + fn) (0 when undefined), weights = support in y_true.
__global__ void metrics_finalize_kernel(const int* __restrict__ confusion, int C, double* __restrict__ sums) {
    const int c = threadIdx.x;
    double tp = 0.0, support = 0.0, predicted = 0.0;
    if (c < C) {
        tp = confusion[c * C + c];
        for (int j = 0; j < C; ++j) {
            support += confusion[c * C + j];
            predicted += confusion[j * C + c];
        }
    }
    const double denom = support + predicted;  // = 2 tp + fp + fn
    double f1w = denom > 0.0 ? (2.0 * tp / denom) * support : 0.0;
    double n = support, hit = tp;
    for (int o = 16; o > 0; o >>= 1) {
        f1w += __shfl_xor_sync(0xffffffffu, f1w, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
        hit += __shfl_xor_sync(0xffffffffu, hit, o);
    }
    if (c == 0) {
        // an empty batch scores 0 for both and still counts as a batch (the reference divides by len(loader))
        sums[0] += n > 0.0 ? hit / n : 0.0;
        sums[1] += n > 0.0 ? f1w / n : 0.0;
        sums[2] += 1.0;
    }
}

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) % 16) == 0; }

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_collate_dialogues(const float* text_table, int64_t text_ld, int32_t Dt, const float* audio_table, int64_t audio_ld,
                                     int32_t Da, const int64_t* labels, const int64_t* utt_rows, const int64_t* dlg_offsets,
                                     const int64_t* batch_dialogues, int64_t B, int64_t S, float* text_out, float* audio_out,
                                     int64_t* emotion_out, uint8_t* mask_out, void* stream) {
    M2F_REQUIRE(text_table && audio_table && labels && utt_rows && dlg_offsets && batch_dialogues && text_out && audio_out && emotion_out &&
                    mask_out,
                M2F_ERR_ARG, "collate_dialogues: null pointer");
    M2F_REQUIRE(B > 0 && S > 0 && Dt > 0 && Da > 0 && text_ld >= Dt && audio_ld >= Da, M2F_ERR_ARG,
                "collate_dialogues: bad extent (B %lld, S %lld, Dt %d, Da %d, ld %lld / %lld)", (long long)B, (long long)S, Dt, Da,
                (long long)text_ld, (long long)audio_ld);
    M2F_REQUIRE(B * S <= 0x7fffffffLL, M2F_ERR_UNSUPPORTED, "collate_dialogues: B * S = %lld exceeds the grid limit", (long long)(B * S));
    const bool vec = (Dt % 4) == 0 && (Da % 4) == 0 && (text_ld % 4) == 0 && (audio_ld % 4) == 0 && al16(text_table) && al16(audio_table) &&
                     al16(text_out) && al16(audio_out);
    auto st = reinterpret_cast<cudaStream_t>(stream);
    auto kern = vec ? collate_kernel<true> : collate_kernel<false>;
    kern<<<(unsigned)(B * S), COLLATE_THREADS, 0, st>>>(text_table, text_ld, Dt, audio_table, audio_ld, Da, labels, utt_rows, dlg_offsets,
                                                        batch_dialogues, (int)S, text_out, audio_out, emotion_out, mask_out);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_batch_metrics(const float* logits, int64_t ld, const int64_t* target, int64_t M, int32_t C, int64_t ignore_index,
                                 int32_t* confusion, double* sums, int64_t* pred, void* stream) {
    M2F_REQUIRE(logits && target && confusion && sums, M2F_ERR_ARG, "batch_metrics: null pointer");
    M2F_REQUIRE(M >= 0 && C > 0 && C <= MET_MAX_C && ld >= C, M2F_ERR_ARG, "batch_metrics: bad extent (M %lld, C %d <= %d, ld %lld)",
                (long long)M, C, MET_MAX_C, (long long)ld);
    auto st = reinterpret_cast<cudaStream_t>(stream);
    M2F_CUDA(cudaMemsetAsync(confusion, 0, sizeof(int32_t) * C * C, st));
    if (M > 0) {
        const int64_t want = (M + MET_THREADS - 1) / MET_THREADS;
        const unsigned grid = (unsigned)(want < 148 * 4 ? want : 148 * 4);
        confusion_kernel<<<grid, MET_THREADS, 0, st>>>(logits, ld, target, M, C, ignore_index, confusion, pred);
        M2F_LAUNCH_CHECK();
    }
    metrics_finalize_kernel<<<1, 32, 0, st>>>(confusion, C, sums);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
