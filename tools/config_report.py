"""Throughput of every BASELINE.json configuration that fits one GPU, in one run, as JSON lines:

  C2  train step (fwd + CE + bwd + Adam), default dims, B=64, S=33, bf16 mode and fp32-parity mode
  C3  train step, text 1024 / audio_mel 300, 2 transformers x 6 layers, 8 fusion layers, 3-layer head, B=64, S=33
  C4  (per-GPU leg) an epoch over a bucketed synthetic split served by the on-device loader: valid utterances / s
  C5  test() body (eval, inference_mode, argmax + on-device metrics), B=512, S=256, bf16 and fp32 mode, full and ragged
  +   m2f_collate_dialogues alone: achieved HBM GB/s (algorithmic bytes = (Dt + Da) * 4 * 2 per valid utterance +
      (Dt + Da) * 4 per padded position)

    python tools/config_report.py [out.json]

Device-timed with CUDA events after warm-up; inputs larger than L2 or rotated between steps are noted per line."""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
D = importlib.import_module("multimodal-emotion-recognition_b200.data")
LINES = []


def emit(**kw):
    LINES.append(kw)
    print(json.dumps(kw), flush=True)


def timed(fn, warmup, steps):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def train_config(name, cfg, planes, B, S, Dt, Da, audio_kind, flop_per_utt):
    torch.manual_seed(0)
    model = pkg.M2FNet(cfg, planes=planes).to("cuda").train()
    crit = pkg.FusedCrossEntropyLoss().cuda()
    opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
    pool = [{k: v.cuda() for k, v in D.synth_batch(B, S, Dt, Da, seed=1234 + i, audio_kind=audio_kind).items()} for i in range(4)]
    it = [0]

    def step():
        b = pool[it[0] % len(pool)]
        it[0] += 1
        opt.zero_grad(set_to_none=True)
        crit(model(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"]).backward()
        opt.step()

    ms = timed(step, 5, 20)
    n_params = sum(p.numel() for p in model.parameters())
    emit(config=name, mode="bf16" if planes == 1 else "fp32-parity (3 bf16 planes)", B=B, S=S, params=n_params, ms_per_step=round(ms, 4),
         utt_per_s=round(B * S / ms * 1e3), tflops=round(3 * flop_per_utt * B * S / ms / 1e9, 1), dropout=cfg.dropout,
         step="fwd + CE + bwd + Adam", inputs="4 device batches rotated")
    del model, opt, pool
    torch.cuda.empty_cache()


def infer_config(name, planes, B, S, ragged):
    torch.manual_seed(0)
    model = pkg.M2FNet(D.default_model_config(), planes=planes).to("cuda").eval()
    b = {k: v.cuda() for k, v in D.synth_batch(B, S, 768, 768, ragged=ragged).items()}
    met = D.BatchMetrics(7, "cuda")

    def step():
        with torch.inference_mode():
            met.update(model(b["text"], b["audio"], b["padding_mask"]), b["emotion"])

    ms = timed(step, 2, 5)
    valid = int((~b["padding_mask"]).sum())
    emit(config=name, mode="bf16" if planes == 1 else "fp32-parity (3 bf16 planes)", B=B, S=S, ragged=ragged, ms_per_batch=round(ms, 3),
         padded_utt_per_s=round(B * S / ms * 1e3), valid_utt_per_s=round(valid / ms * 1e3), tflops=round(185.6e6 * B * S / ms / 1e9, 1),
         step="test() body: forward, argmax, confusion matrix, accuracy / weighted F1 on device", metrics=met.result(),
         workspace_gb=round(sum(p.ws_bytes for p in model._plans.values()) / 1e9, 2), inputs="805 MB per batch (> L2)")
    del model, b
    torch.cuda.empty_cache()


def synthetic_store(n_dlg, Dt, Da, seed=0):
    g = np.random.default_rng(seed)
    lengths = np.clip(np.round(np.exp(g.normal(2.1, 0.6, n_dlg))), 1, 33).astype(np.int64)
    did = np.repeat(np.arange(n_dlg), lengths)
    uid = np.concatenate([np.arange(n) for n in lengths])
    n = int(lengths.sum())
    gen = torch.Generator(device="cuda").manual_seed(seed)
    text = torch.randn(n, Dt, device="cuda", generator=gen) * 0.63 + 0.025
    audio = torch.randn(n, Da, device="cuda", generator=gen) * 0.233
    labels = torch.multinomial(torch.tensor(D.MELD_CLASS_P), n, replacement=True)
    return D.DeviceDialogueStore(text, audio, labels, did, uid, "cuda"), n


def collate_bandwidth():
    store, n = synthetic_store(20000, 768, 768)
    for label, idx, S in (("full-length rows", np.nonzero(store.lengths >= 24)[0][:2048], 33), ("MELD-like lengths", np.arange(8192), 33)):
        live = int(store.lengths[idx].sum())
        pad = len(idx) * S - live
        byts = (768 + 768) * 4 * (2 * live + pad) + len(idx) * S * 9
        ms = timed(lambda: store.collate(idx, S), 3, 20)
        emit(config="collate", case=label, B=len(idx), S=S, valid=live, ms=round(ms, 4), algorithmic_mb=round(byts / 1e6, 1),
             gb_per_s=round(byts / ms / 1e6, 1), note="includes the H2D copy of the B dialogue indices and 4 torch.empty calls")
    return store, n


def bucketed_epoch(store, n_utt):
    """C4's per-GPU leg with the real shape of the data: MELD-like lengths, bucketed batches, train step per batch."""
    torch.manual_seed(0)
    model = pkg.M2FNet(D.default_model_config(), planes=1).to("cuda").train()
    crit = pkg.FusedCrossEntropyLoss().cuda()
    opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
    for buckets in ((33,), (8, 16, 24, 33)):
        sampler = D.LengthBucketSampler(store.lengths, 64, buckets=buckets, seed=1, drop_last=True)
        loader = D.DeviceLoader(store, sampler)

        def epoch():
            for b in loader:
                opt.zero_grad(set_to_none=True)
                crit(model(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"]).backward()
                opt.step()

        ms = timed(epoch, 1, 1)
        live = sum(int(store.lengths[i].sum()) for i, _ in sampler)
        emit(config="C4 per-GPU leg (20k dialogues, MELD-like lengths, B=64)", buckets=list(buckets), batches=len(sampler),
             padded_fraction=round(sampler.padded_fraction(), 3), epoch_ms=round(ms, 1), valid_utt_per_s=round(live / ms * 1e3),
             step="collate on device + fwd + CE + bwd + Adam", mode="bf16")


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else None
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    base = D.default_model_config()
    train_config("C2", base, 1, 64, 33, 768, 768, "wav2vec2", 174.0e6)
    train_config("C2", base, 3, 64, 33, 768, 768, "wav2vec2", 174.0e6)
    c3 = D.default_model_config(TEXT=dict(embedding_size=1024, n_head=16, n_transformers=2, n_encoder_layers=6),
                                AUDIO=dict(embedding_size=300, n_head=10, n_transformers=2, n_encoder_layers=6),
                                FAM=dict(embedding_size=1024, n_head=16, n_layers=8), CLASSIFIER=dict(hidden_size=1024, output_size=7, n_layers=3))
    train_config("C3", c3, 1, 64, 33, 1024, 300, "mel", 352.3e6)
    train_config("C3", c3, 3, 64, 33, 1024, 300, "mel", 352.3e6)
    store, n = collate_bandwidth()
    bucketed_epoch(store, n)
    del store
    torch.cuda.empty_cache()
    infer_config("C5", 1, 512, 256, False)
    infer_config("C5", 1, 512, 256, True)
    infer_config("C5", 3, 512, 256, False)
    if out:
        json.dump(LINES, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
