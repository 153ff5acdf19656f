"""GPU parity of the rows either side of the model step, through the C ABI:
m2f_collate_dialogues against the reference's recorded batches and the collate oracle (bit-exact: copies and
constants), m2f_batch_metrics against sklearn on the host exactly as src/test.py:65-74 calls it."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import collate_oracle as CO

from .conftest import PKG_NAME

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def D(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    return importlib.import_module(PKG_NAME + ".data")


def _same(got, ref):
    for k in ("text", "audio", "padding_mask", "emotion"):
        g = got[k].cpu()
        r = torch.as_tensor(ref[k])
        assert g.dtype == r.dtype and g.shape == r.shape, k
        assert torch.equal(g, r), k


def test_collate_reproduces_the_reference_batches(D, golden_dir):
    z = np.load(os.path.join(golden_dir, "collate_ref.npz"))
    store = D.DeviceDialogueStore(torch.from_numpy(z["text_emb"]), torch.from_numpy(z["audio_emb"]), z["labels"], z["dialogue_id"],
                                  z["utterance_id"], "cuda")
    assert len(store) == len(z["dataset_dialogue_ids"])
    for bi in range(int(z["n_batches"])):
        got = store.collate(z[f"batch{bi}/indices"])  # Da = 6: the scalar path
        _same(got, {k: z[f"batch{bi}/{k}"] for k in ("text", "audio", "padding_mask", "emotion")})
    longest = int(np.argmax(store.lengths))
    with pytest.raises(ValueError, match="shorter than the longest"):
        store.collate([0, longest], S=int(store.lengths[longest]) - 1)
    with pytest.raises(IndexError):
        store.collate([7])


@pytest.mark.parametrize("Dt,Da", [(768, 768), (1024, 300), (10, 7)])
def test_collate_matches_oracle_on_a_shuffled_split(D, Dt, Da):
    g = np.random.default_rng(11)
    n_dlg = 40
    lengths = np.clip(np.round(np.exp(g.normal(2.1, 0.6, n_dlg))), 1, 33).astype(int)
    lengths[0], lengths[1] = 33, 1
    ids = g.permutation(500)[:n_dlg]
    rows = [(int(ids[d]), u) for d in range(n_dlg) for u in range(lengths[d])]
    rows = [rows[i] for i in g.permutation(len(rows))]  # CSV order: dialogues interleaved, utterances out of order
    did = np.array([r[0] for r in rows])
    uid = np.array([r[1] for r in rows])
    labels = g.integers(0, 7, len(rows))
    text = torch.from_numpy(g.standard_normal((len(rows), Dt)).astype(np.float32))
    audio = torch.from_numpy(g.standard_normal((len(rows), Da)).astype(np.float32))
    store = D.DeviceDialogueStore(text, audio, labels, did, uid, "cuda")
    for idx, S in (([0, 1, 2, 3, 4, 5, 6, 7], None), ([1], None), ([5, 9, 9, 30], 33), (list(range(n_dlg)), 40)):
        _same(store.collate(idx, S), CO.collate(idx, did, uid, labels, text, audio, S))
    sampler = D.LengthBucketSampler(store.lengths, 8, seed=2)
    seen = 0
    for batch, (idx, S) in zip(D.DeviceLoader(store, sampler), sampler):
        assert batch["text"].shape == (len(idx), S, Dt) and batch["text"].is_cuda
        _same(batch, CO.collate(idx, did, uid, labels, text, audio, S))
        seen += int((~batch["padding_mask"]).sum())
    assert seen == len(rows)


def test_batch_metrics_match_sklearn_over_an_epoch(D):
    from sklearn.metrics import accuracy_score, f1_score
    g = torch.Generator().manual_seed(21)
    met = D.BatchMetrics(7, "cuda")
    acc = f1 = 0.0
    n_batches = 5
    for i in range(n_batches):
        B, S = (64, 33) if i else (512, 256)
        logits = torch.randn(B, S, 7, generator=g)
        if i == 1:
            logits[..., 4:] = -50.0  # classes 4..6 never predicted
        if i == 2:
            logits = torch.round(logits)  # ties: the first maximum wins, as torch.argmax
        emotion = torch.randint(0, 7 if i != 3 else 3, (B, S), generator=g)
        L = torch.randint(1, S + 1, (B,), generator=g)
        emotion[torch.arange(S)[None, :] >= L[:, None]] = -1
        pred = met.update(logits.cuda(), emotion.cuda(), want_pred=True)
        assert torch.equal(pred.cpu(), logits.argmax(dim=2))
        m = emotion != -1  # src/test.py:66-71
        y, yp = emotion[m].flatten().numpy(), logits.argmax(dim=2)[m].flatten().numpy()
        acc += accuracy_score(y, yp)
        f1 += f1_score(y, yp, average="weighted")
        a_b, f_b = D.scores_from_confusion(met.confusion.cpu())
        assert abs(a_b - accuracy_score(y, yp)) < 1e-12 and abs(f_b - f1_score(y, yp, average="weighted")) < 1e-12
    got = met.result()
    assert abs(got[0] - acc / n_batches) < 1e-12 and abs(got[1] - f1 / n_batches) < 1e-12
    met.reset()
    assert met.result() == (0.0, 0.0)
    with pytest.raises(ValueError, match="do not match"):
        met.update(torch.zeros(2, 3, 5, device="cuda"), torch.zeros(2, 3, dtype=torch.long, device="cuda"))


def test_test_loop_body_on_device_equals_the_host_path(D, pkg):
    """The reference's test() (src/test.py:51-74) over a bucketed DeviceLoader with on-device metrics, against the same
    loop scoring each batch with sklearn on the host."""
    from sklearn.metrics import accuracy_score, f1_score
    from oracle import m2fnet_oracle as O
    cfg = O.default_model_cfg(dropout=0.4, TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=1),
                              AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=1), FAM=dict(embedding_size=64, n_head=4, n_layers=1),
                              CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2))
    model = pkg.M2FNet(cfg, planes=3)
    model.load_state_dict({k: v.float() for k, v in O.synth_params(cfg).items()}, strict=True)
    model = model.to("cuda").eval()
    g = np.random.default_rng(4)
    lengths = np.clip(np.round(np.exp(g.normal(2.1, 0.6, 60))), 1, 33).astype(int)
    did = np.repeat(np.arange(60), lengths)
    uid = np.concatenate([np.arange(n) for n in lengths])
    labels = g.integers(0, 7, did.size)
    store = D.DeviceDialogueStore(torch.from_numpy(g.standard_normal((did.size, 128)).astype(np.float32)),
                                  torch.from_numpy(g.standard_normal((did.size, 64)).astype(np.float32) * 0.2), labels, did, uid, "cuda")
    loader = D.DeviceLoader(store, D.LengthBucketSampler(store.lengths, 16, shuffle=False))
    met = D.BatchMetrics(7, "cuda")
    acc = f1 = 0.0
    with torch.inference_mode():
        for data in loader:
            outputs = model(data["text"], data["audio"], data["padding_mask"])
            met.update(outputs, data["emotion"])
            pred = torch.argmax(outputs, dim=2)
            mask = data["emotion"] != -1
            y, yp = data["emotion"][mask].flatten().cpu().numpy(), pred[mask].flatten().cpu().numpy()
            acc += accuracy_score(y, yp)
            f1 += f1_score(y, yp, average="weighted")
    got = met.result()
    assert abs(got[0] - acc / len(loader)) < 1e-12 and abs(got[1] - f1 / len(loader)) < 1e-12
