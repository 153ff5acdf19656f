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
        # every rank must run the SAME number of steps (the gradient exchange is collective: a rank with one batch more
        # would wait for peers that never come): all ranks keep the size of the smallest shard, i.e. at most world - 1
        # dialogues of the epoch are dropped (the DistributedSampler(drop_last=True) convention)
        self.hi = self.lo + n_dialogues // world
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


def scores_from_confusion(cm) -> tuple:
    """(accuracy, weighted F1) of one batch from its C x C confusion matrix (row = true class, column = predicted), as
    sklearn computes them for src/test.py:70-71: F1 per class present in y_true or y_pred, 0 where undefined, weighted
    by the support in y_true.  Host arithmetic on C*C counts; the device computes the same in m2f_batch_metrics."""
    cm = torch.as_tensor(cm).double()
    n = cm.sum().item()
    if n == 0:
        return 0.0, 0.0
    tp, support, predicted = cm.diag(), cm.sum(1), cm.sum(0)
    denom = support + predicted
    f1 = torch.where(denom > 0, 2 * tp / denom.clamp_min(1), torch.zeros_like(tp))
    return tp.sum().item() / n, (f1 * support).sum().item() / n


class BatchMetrics:
    """Running accuracy / weighted-F1 of the reference's validate() and test() loops (src/train.py:262-267,
    src/test.py:65-74: per-batch sklearn scores, averaged over the batches) kept on the device: `update` launches
    m2f_batch_metrics on the current stream (argmax, confusion matrix, both scores added to three doubles) and never
    synchronises; `result()` is the one device -> host read of the epoch."""

    def __init__(self, num_classes: int, device, ignore_index: int = -1):
        self.C, self.ignore_index = int(num_classes), int(ignore_index)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("BatchMetrics runs on a CUDA device only (no CPU fallback)")
        self.confusion = torch.zeros(self.C, self.C, dtype=torch.int32, device=self.device)  # of the last batch
        self.sums = torch.zeros(3, dtype=torch.float64, device=self.device)

    def reset(self):
        self.sums.zero_()

    @torch.no_grad()
    def update(self, outputs: torch.Tensor, emotion: torch.Tensor, want_pred: bool = False):
        """outputs [B,S,C] fp32 logits, emotion [B,S] int64 with `ignore_index` at padded positions."""
        from . import _lib
        if not (outputs.is_cuda and emotion.is_cuda):
            raise RuntimeError("BatchMetrics.update: CUDA tensors required")
        if outputs.shape[-1] != self.C or outputs.shape[:-1] != emotion.shape:
            raise ValueError(f"BatchMetrics.update: logits {tuple(outputs.shape)} do not match targets {tuple(emotion.shape)} x {self.C} classes")
        lg = outputs if (outputs.dtype == torch.float32 and outputs.is_contiguous()) else outputs.float().contiguous()
        tg = emotion if (emotion.dtype == torch.int64 and emotion.is_contiguous()) else emotion.long().contiguous()
        pred = torch.empty(emotion.shape, dtype=torch.int64, device=outputs.device) if want_pred else None
        L = _lib.lib()
        _lib.check(L.m2f_batch_metrics(lg.data_ptr(), self.C, tg.data_ptr(), tg.numel(), self.C, self.ignore_index, self.confusion.data_ptr(),
                                       self.sums.data_ptr(), pred.data_ptr() if want_pred else None,
                                       torch.cuda.current_stream(outputs.device).cuda_stream), "batch_metrics")
        return pred

    def result(self) -> tuple:
        """(accuracy, weighted_f1) averaged over the batches seen since reset(), as test() returns them."""
        acc, f1, n = self.sums.tolist()
        return (acc / n, f1 / n) if n else (0.0, 0.0)


class DialogueIndex:
    """Host-side index of a MELD split in the order the reference Dataset uses (src/dataset.py:26-35): dialogues in
    order of first appearance of their Dialogue_ID (`unique()`), each dialogue's table rows sorted by Utterance_ID.
    `rows[offsets[d]:offsets[d+1]]` are the CSV / embedding-table rows of dialogue d."""

    def __init__(self, dialogue_id, utterance_id):
        import numpy as np
        did = np.asarray(dialogue_id).astype(np.int64).ravel()
        uid = np.asarray(utterance_id).astype(np.int64).ravel()
        if did.shape != uid.shape:
            raise ValueError("DialogueIndex: Dialogue_ID and Utterance_ID columns differ in length")
        uniq, first = np.unique(did, return_index=True)
        order = uniq[np.argsort(first, kind="stable")]          # first-appearance order, as pandas unique()
        rank = np.empty(int(did.max()) + 1 if did.size else 0, dtype=np.int64)
        rank[order] = np.arange(order.size)
        key = rank[did]
        perm = np.lexsort((uid, key))                            # by dialogue, then Utterance_ID
        self.dialogue_ids = order
        self.rows = perm.astype(np.int64)
        self.lengths = np.bincount(key, minlength=order.size).astype(np.int64)
        self.offsets = np.concatenate(([0], np.cumsum(self.lengths))).astype(np.int64)
        k2, u2 = key[perm], uid[perm]
        if np.any((k2[1:] == k2[:-1]) & (u2[1:] == u2[:-1])):    # the reference asserts one row per (dialogue, utterance)
            raise ValueError("DialogueIndex: duplicate (Dialogue_ID, Utterance_ID) pair")

    def __len__(self):
        return int(self.dialogue_ids.size)


class DeviceDialogueStore:
    """The embedding tables of one split resident in HBM, batches assembled by m2f_collate_dialogues instead of
    Dataset.__getitem__ (a DataFrame scan per utterance) + collate_fn.  `text_emb` / `audio_emb`
    are the [n_rows, D] tensors of the reference's pickles, `labels` the mapped Emotion column (one per CSV row).
    MELD train is 9989 x (768 + 768) fp32 = 61 MB; a 100 k-dialogue synthetic epoch at S = 33 is 20 GB."""

    def __init__(self, text_emb: torch.Tensor, audio_emb: torch.Tensor, labels, dialogue_id, utterance_id, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DeviceDialogueStore needs a CUDA device (no CPU fallback)")
        self.index = DialogueIndex(dialogue_id, utterance_id)
        n = self.index.rows.size
        if text_emb.shape[0] != n or audio_emb.shape[0] != n or len(labels) != n:
            raise ValueError("DeviceDialogueStore: tables, labels and id columns must have one entry per CSV row")
        self.text = text_emb.to(self.device, torch.float32).contiguous()
        self.audio = audio_emb.to(self.device, torch.float32).contiguous()
        self.labels = torch.as_tensor(labels).to(self.device, torch.int64).contiguous()
        self.rows = torch.from_numpy(self.index.rows).to(self.device)
        self.offsets = torch.from_numpy(self.index.offsets).to(self.device)
        self.lengths = self.index.lengths

    def __len__(self):
        return len(self.index)

    def collate(self, dialogues, S: Optional[int] = None) -> Dict[str, torch.Tensor]:
        """The dict collate_fn returns (src/dataset.py:71-89), as device tensors.  `dialogues`: indices into the
        dataset (what the DataLoader's sampler would hand to __getitem__); S defaults to the in-batch maximum length
        (apply_padding, src/utils.py:27), a larger S pads further (static shapes for plans / graphs)."""
        import numpy as np
        from . import _lib
        idx = np.asarray(dialogues, dtype=np.int64).ravel()
        if idx.size == 0:
            raise ValueError("collate: empty batch")
        if idx.min() < 0 or idx.max() >= len(self):
            raise IndexError("collate: dialogue index out of range")
        longest = int(self.lengths[idx].max())
        S = longest if S is None else int(S)
        if S < longest:
            raise ValueError(f"collate: S = {S} is shorter than the longest selected dialogue ({longest})")
        B, Dt, Da = int(idx.size), self.text.shape[1], self.audio.shape[1]
        batch = torch.from_numpy(idx).to(self.device, non_blocking=True)
        text = torch.empty(B, S, Dt, device=self.device)
        audio = torch.empty(B, S, Da, device=self.device)
        emotion = torch.empty(B, S, dtype=torch.int64, device=self.device)
        mask = torch.empty(B, S, dtype=torch.bool, device=self.device)
        _lib.check(_lib.lib().m2f_collate_dialogues(self.text.data_ptr(), Dt, Dt, self.audio.data_ptr(), Da, Da, self.labels.data_ptr(),
                                                    self.rows.data_ptr(), self.offsets.data_ptr(), batch.data_ptr(), B, S, text.data_ptr(),
                                                    audio.data_ptr(), emotion.data_ptr(), mask.data_ptr(),
                                                    torch.cuda.current_stream(self.device).cuda_stream), "collate_dialogues")
        return {"text": text, "audio": audio, "padding_mask": mask, "emotion": emotion}


class LengthBucketSampler:
    """Batches of dialogue indices grouped by length so that a batch pads to its bucket's S instead of to one long
    outlier (MELD: mean 9-10 utterances, maximum 33; in-batch-max padding of a random batch wastes most positions).
    `buckets` are the padded lengths on offer (ascending; the last must cover the longest dialogue): every batch comes
    from one bucket and is reported as (indices, S_bucket), so the model sees a handful of static (B, S) shapes - one
    plan and one CUDA graph each.  Deterministic in (seed, epoch); shards by rank like parallel.shard_range does for
    the contiguous case: every rank draws the same batches and keeps those with index % world == rank."""

    def __init__(self, lengths, batch_size: int, buckets=(8, 16, 24, 33), shuffle: bool = True, seed: int = 0, rank: int = 0,
                 world: int = 1, drop_last: bool = False):
        import numpy as np
        self.lengths = np.asarray(lengths, dtype=np.int64)
        self.buckets = tuple(sorted(int(b) for b in buckets))
        if self.lengths.size and int(self.lengths.max()) > self.buckets[-1]:
            raise ValueError(f"LengthBucketSampler: longest dialogue ({int(self.lengths.max())}) exceeds the last bucket ({self.buckets[-1]})")
        self.bs, self.shuffle, self.seed, self.rank, self.world, self.drop_last = int(batch_size), shuffle, int(seed), rank, world, drop_last
        self.epoch = 0
        self.which = np.searchsorted(np.asarray(self.buckets), self.lengths, side="left")  # smallest bucket >= length

    def set_epoch(self, epoch: int):
        self.epoch = int(epoch)

    def _batches(self):
        import numpy as np
        rng = np.random.default_rng(self.seed + 7919 * self.epoch)
        out = []
        for bi, S in enumerate(self.buckets):
            members = np.nonzero(self.which == bi)[0]
            if self.shuffle:
                members = rng.permutation(members)
            for i in range(0, members.size, self.bs):
                chunk = members[i:i + self.bs]
                if chunk.size < self.bs and self.drop_last:
                    continue
                out.append((chunk, S))
        if self.shuffle:
            out = [out[i] for i in rng.permutation(len(out))]
        # equal step counts on every rank (the gradient exchange is collective): the last len(out) % world batches of the
        # epoch's (reshuffled every epoch) order are dropped, DistributedSampler(drop_last=True) style
        out = out[:len(out) // self.world * self.world]
        return out[self.rank::self.world]

    def __len__(self):
        return len(self._batches())

    def __iter__(self):
        return iter(self._batches())

    def padded_fraction(self) -> float:
        """Share of the (B, S) positions this sampler's batches spend on padding (all ranks)."""
        rank, world = self.rank, self.world
        self.rank, self.world = 0, 1
        try:
            tot = sum(len(c) * S for c, S in self._batches())
            live = sum(int(self.lengths[c].sum()) for c, _ in self._batches())
        finally:
            self.rank, self.world = rank, world
        return 1.0 - live / tot if tot else 0.0


class DeviceLoader:
    """Iterable the reference's train() / validate() / test() loops accept in place of their DataLoader
    (src/train.py:222-227, src/test.py:58-62): yields collate_fn's dict already on the device (`.to(device)` is then
    a no-op), batches chosen by `sampler` (LengthBucketSampler, or any iterable of index lists / (indices, S))."""

    def __init__(self, store: DeviceDialogueStore, sampler):
        self.store, self.sampler = store, sampler

    def __len__(self):
        return len(self.sampler)

    def __iter__(self):
        for item in self.sampler:
            idx, S = item if (isinstance(item, tuple) and len(item) == 2) else (item, None)
            yield self.store.collate(idx, S)


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
