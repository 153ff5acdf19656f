"""CPU side of the batch-assembly / metrics rows: the collate oracle is pinned against what the
reference's own Dataset + collate_fn produced (tests/golden/collate_ref.npz, written by oracle/gen_golden_collate.py),
and the host logic around the kernels (dialogue index, length buckets, score arithmetic) is checked against it."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import collate_oracle as CO

from .conftest import PKG_NAME


@pytest.fixture(scope="module")
def Z(golden_dir):
    z = np.load(os.path.join(golden_dir, "collate_ref.npz"))
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="module")
def D(pkg):
    return importlib.import_module(PKG_NAME + ".data")


def test_collate_oracle_reproduces_the_reference_batches(Z):
    for bi in range(int(Z["n_batches"])):
        got = CO.collate(Z[f"batch{bi}/indices"], Z["dialogue_id"], Z["utterance_id"], Z["labels"], Z["text_emb"], Z["audio_emb"])
        for k in ("text", "audio", "padding_mask", "emotion"):
            ref = Z[f"batch{bi}/{k}"]
            assert got[k].shape == ref.shape and got[k].numpy().dtype == ref.dtype, (bi, k)
            assert np.array_equal(got[k].numpy(), ref), (bi, k)  # bit-exact: copies and constants only


def test_dialogue_index_orders_like_the_reference_dataset(Z, D):
    ix = D.DialogueIndex(Z["dialogue_id"], Z["utterance_id"])
    assert np.array_equal(ix.dialogue_ids, Z["dataset_dialogue_ids"])  # Dataset.dialogue_ids (src/dataset.py:26)
    for d in range(len(ix)):
        rows = ix.rows[ix.offsets[d]:ix.offsets[d + 1]]
        assert (Z["dialogue_id"][rows] == ix.dialogue_ids[d]).all()
        assert np.array_equal(Z["utterance_id"][rows], np.sort(Z["utterance_id"][rows]))
        _, _, emo = CO.getitem(d, Z["dialogue_id"], Z["utterance_id"], Z["labels"], Z["text_emb"], Z["audio_emb"])
        assert np.array_equal(Z["labels"][rows], emo.numpy())
    with pytest.raises(ValueError, match="duplicate"):
        D.DialogueIndex([1, 1, 2], [0, 0, 0])


def test_length_buckets_cover_every_dialogue_once(D):
    g = np.random.default_rng(0)
    lengths = np.clip(np.round(np.exp(g.normal(2.1, 0.6, 1000))), 1, 33).astype(int)
    world = 3
    seen = []
    for rank in range(world):
        s = D.LengthBucketSampler(lengths, 16, buckets=(8, 16, 24, 33), seed=5, rank=rank, world=world)
        s.set_epoch(2)
        again = D.LengthBucketSampler(lengths, 16, buckets=(8, 16, 24, 33), seed=5, rank=rank, world=world)
        again.set_epoch(2)
        for (idx, S), (idx2, S2) in zip(s, again):
            assert S == S2 and np.array_equal(idx, idx2)  # deterministic in (seed, epoch)
            assert lengths[idx].max() <= S and len(idx) <= 16
            assert S == 8 or lengths[idx].min() > {16: 8, 24: 16, 33: 24}[S]  # smallest bucket that fits
            seen.extend(idx.tolist())
    # every rank gets the same number of batches (the gradient exchange is collective), so up to world - 1 batches of the
    # epoch's order are dropped; nothing is served twice
    assert len(seen) == len(set(seen)) and len(seen) >= 1000 - 16 * (world - 1)
    s = D.LengthBucketSampler(lengths, 16, buckets=(8, 16, 24, 33), seed=5)
    assert sorted(i for idx, _ in s for i in idx.tolist()) == list(range(1000))  # one rank: every dialogue exactly once
    one_bucket = D.LengthBucketSampler(lengths, 16, buckets=(33,), seed=5)
    assert s.padded_fraction() < 0.35 < one_bucket.padded_fraction()  # MELD-like lengths: most of a 33-wide batch is padding
    with pytest.raises(ValueError, match="exceeds the last bucket"):
        D.LengthBucketSampler([40], 4, buckets=(8, 33))


def test_scores_from_confusion_equal_sklearn(D):
    from sklearn.metrics import accuracy_score, confusion_matrix, f1_score
    g = np.random.default_rng(3)
    for trial in range(6):
        C = 7
        y = g.integers(0, C if trial % 2 else 4, 200)  # odd trials: every class may occur; even: classes 4..6 never true
        yp = g.integers(0, C if trial < 4 else 3, 200)
        cm = confusion_matrix(y, yp, labels=list(range(C)))
        acc, f1 = D.scores_from_confusion(cm)
        assert abs(acc - accuracy_score(y, yp)) < 1e-12
        assert abs(f1 - f1_score(y, yp, average="weighted")) < 1e-12
    assert D.scores_from_confusion(np.zeros((7, 7))) == (0.0, 0.0)


def test_device_only_classes_refuse_the_cpu(D):
    with pytest.raises(RuntimeError, match="CUDA"):
        D.BatchMetrics(7, "cpu")
    with pytest.raises(RuntimeError, match="CUDA"):
        D.DeviceDialogueStore(torch.zeros(2, 4), torch.zeros(2, 4), [0, 1], [0, 0], [0, 1], "cpu")
