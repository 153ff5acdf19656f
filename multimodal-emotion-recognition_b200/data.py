"""Input side of the hot path: config objects with attribute access (what `Munch` gives the reference,
src/utils.py:8-13), MELD-shaped synthetic batches laid out exactly as the reference's `collate_fn`
does (src/dataset.py:71-89: text/audio zero-padded to the in-batch max, padding_mask True at padding, emotion -1 at
padding), an in-memory loader yielding them, and the per-batch metrics of src/test.py:65-74 computed on the device."""
from __future__ import annotations

from typing import Dict, Iterator, Optional

import torch


class Config(dict):
    """dict with attribute access, recursively (a stand-in for munch.Munch)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    @classmethod
    def wrap(cls, d):
        if isinstance(d, dict):
            return cls({k: cls.wrap(v) for k, v in d.items()})
        return d

    @classmethod
    def fromYAML(cls, stream):
        import yaml
        return cls.wrap(yaml.safe_load(stream))


def default_model_config(**over) -> Config:
    """The `model:` block of the reference's src/config.yaml:31-56 (defaults)."""
    c = {
        "dropout": 0.4,
        "AUDIO": {"enabled": True, "embedding_size": 768, "n_head": 8, "n_transformers": 1, "n_encoder_layers": 6},
        "TEXT": {"enabled": True, "embedding_size": 768, "n_head": 8, "n_transformers": 1, "n_encoder_layers": 6},
        "FAM": {"enabled": True, "embedding_size": 768, "n_head": 8, "n_layers": 5},
        "CLASSIFIER": {"hidden_size": 768, "output_size": 7, "n_layers": 2},
    }
    for k, v in over.items():
        if isinstance(v, dict):
            c[k].update(v)
        else:
            c[k] = v
    return Config.wrap(c)


MELD_CLASS_P = (.47, .17, .07, .11, .12, .03, .03)  # neutral, joy, sadness, anger, surprise, fear, disgust (src/dataset.py:22)


def synth_batch(B: int, S: int, Dt: int, Da: int, C: int = 7, seed: int = 1234, ragged: bool = False,
                audio_kind: str = "wav2vec2") -> Dict[str, torch.Tensor]:
    """One collated batch (CPU tensors) with the value statistics of the shipped embeddings."""
    g = torch.Generator().manual_seed(seed)
    text = torch.randn(B, S, Dt, generator=g) * 0.63 + 0.025
    if audio_kind == "mel":
        audio = torch.randn(B, S, Da, generator=g)
        audio = audio / audio.norm(dim=-1, keepdim=True)
    else:
        audio = torch.randn(B, S, Da, generator=g) * 0.233
    probs = torch.tensor(MELD_CLASS_P)[:C]
    emotion = torch.multinomial(probs / probs.sum(), B * S, replacement=True, generator=g).reshape(B, S)
    if ragged:
        L = torch.exp(torch.randn(B, generator=g) * 0.6 + 2.1).round().clamp(1, S).long()
        L[0] = S
        if B > 1:
            L[B - 1] = 1
    else:
        L = torch.full((B,), S, dtype=torch.long)
    mask = torch.arange(S)[None, :] >= L[:, None]
    return {"text": text.masked_fill(mask[..., None], 0.0), "audio": audio.masked_fill(mask[..., None], 0.0),
            "padding_mask": mask, "emotion": emotion.masked_fill(mask, -1)}


class SyntheticMELD:
    """In-memory loader over `n_dialogues` synthetic dialogues, sharded contiguously by dialogue for data parallelism
    (parallel.shard_range).  Yields the dict `collate_fn` yields; batches are generated on the fly from (seed, index),
    so a 100k-dialogue epoch (BASELINE.json configs[3]) needs no storage."""

    def __init__(self, n_dialogues: int, batch_size: int, S: int = 33, Dt: int = 768, Da: int = 768, C: int = 7, seed: int = 1234,
                 ragged: bool = False, rank: int = 0, world: int = 1, audio_kind: str = "wav2vec2", pin: bool = False,
                 pool: int = 0):
        from .parallel import shard_range
        self.lo, self.hi = shard_range(n_dialogues, rank, world)
        self.bs, self.S, self.Dt, self.Da, self.C = batch_size, S, Dt, Da, C
        self.seed, self.ragged, self.audio_kind, self.pin = seed, ragged, audio_kind, pin
        self.pool = [self._make(i) for i in range(pool)] if pool else None

    def _make(self, i: int, n: Optional[int] = None):
        b = synth_batch(n or self.bs, self.S, self.Dt, self.Da, self.C, seed=self.seed + 7919 * (self.lo + i), ragged=self.ragged,
                        audio_kind=self.audio_kind)
        return {k: v.pin_memory() for k, v in b.items()} if self.pin else b

    def __len__(self):
        return (self.hi - self.lo + self.bs - 1) // self.bs

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        n = self.hi - self.lo
        for i in range(len(self)):
            cnt = min(self.bs, n - i * self.bs)
            if self.pool and cnt == self.bs:
                yield self.pool[i % len(self.pool)]
            else:
                yield self._make(i, cnt)


@torch.no_grad()
def batch_metrics_device(logits: torch.Tensor, emotion: torch.Tensor):
    """Accuracy and weighted-F1 of ONE batch, as src/test.py:65-71 computes them with sklearn on the host (F1 per class
    present in y_true or y_pred, weighted by the support in y_true), from an on-device confusion matrix: one small D2H
    copy of C*C counts instead of the per-batch `.cpu().numpy()` of predictions and labels."""
    C = logits.shape[-1]
    pred = logits.argmax(dim=-1)
    m = emotion != -1
    y, yp = emotion[m], pred[m]
    cm = torch.bincount(y * C + yp, minlength=C * C).reshape(C, C).cpu().double()
    n = cm.sum().item()
    tp = cm.diag()
    support, predicted = cm.sum(1), cm.sum(0)
    denom = support + predicted
    f1 = torch.where(denom > 0, 2 * tp / denom.clamp_min(1), torch.zeros_like(tp))
    present = denom > 0
    acc = tp.sum().item() / n if n else 0.0
    f1w = (f1 * support)[present].sum().item() / n if n else 0.0
    return acc, f1w


class DevicePrefetcher:
    """Host -> device staging for the reference's train()/test() loops (src/train.py:223-227, src/test.py:59-62: one
    `.to(device)` per tensor per batch): wraps any iterable of batch dicts, keeps the host copies in pinned memory and
    issues the copy of batch i+1 on its own stream while batch i computes.  Iterating yields device batches whose
    copies the current stream has been ordered after."""

    def __init__(self, loader, device, depth: int = 2):
        self.loader, self.device, self.depth = loader, torch.device(device), max(1, int(depth))
        self.stream = torch.cuda.Stream(device=self.device)

    def _stage(self, batch):
        with torch.cuda.stream(self.stream):
            dev = {k: (v if v.is_pinned() else v.pin_memory()).to(self.device, non_blocking=True) for k, v in batch.items()}
        ev = torch.cuda.Event()
        ev.record(self.stream)
        return dev, ev

    def __iter__(self):
        it = iter(self.loader)
        queue = []
        try:
            while len(queue) < self.depth:
                queue.append(self._stage(next(it)))
        except StopIteration:
            pass
        while queue:
            dev, ev = queue.pop(0)
            torch.cuda.current_stream(self.device).wait_event(ev)
            for v in dev.values():  # the caching allocator must not hand the memory out while the consumer still reads it
                v.record_stream(torch.cuda.current_stream(self.device))
            try:
                queue.append(self._stage(next(it)))
            except StopIteration:
                pass
            yield dev
