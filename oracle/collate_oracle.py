"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's batch assembly, the checker for
m2f_collate_dialogues / DeviceDialogueStore (never imported by the product package).

Follows, step by step:
  * Dataset.__init__ / __getitem__ (src/dataset.py:26-65): dialogues in order of first appearance of Dialogue_ID
    (`unique()`), the utterances of one dialogue sorted by Utterance_ID, each fetching row `df_row_idx` of the text /
    audio embedding tables and its mapped Emotion label;
  * collate_fn (src/dataset.py:71-89) + apply_padding (src/utils.py:15-31): pad every dialogue of the batch to the
    in-batch maximum length with 0 (features) / -1 (labels); padding_mask = (emotion == -1).

Pinned by tests/test_oracle_golden.py against tests/golden/collate_ref.npz, which oracle/gen_golden_collate.py wrote
by running the reference's own Dataset + collate_fn on a small synthetic split.
"""
import numpy as np
import torch


def getitem(idx, dialogue_id, utterance_id, labels, text_emb, audio_emb):
    """Dataset.__getitem__(idx) (src/dataset.py:32-65) as plain loops."""
    did = np.asarray(dialogue_id)
    uid = np.asarray(utterance_id)
    seen = []
    for d in did.tolist():  # pandas unique(): order of first appearance (src/dataset.py:26)
        if d not in seen:
            seen.append(d)
    dlg = seen[idx]
    rows = [r for r in range(len(did)) if did[r] == dlg]
    rows.sort(key=lambda r: uid[r])  # sort_values(by="Utterance_ID") (src/dataset.py:35)
    text = torch.stack([torch.as_tensor(text_emb[r]) for r in rows])
    audio = torch.stack([torch.as_tensor(audio_emb[r]) for r in rows])
    emotion = torch.tensor([int(labels[r]) for r in rows], dtype=torch.int64)
    return text, audio, emotion


def collate(indices, dialogue_id, utterance_id, labels, text_emb, audio_emb, S=None):
    """collate_fn over [getitem(i) for i in indices] (src/dataset.py:71-89).  S = None pads to the in-batch maximum as
    apply_padding does (src/utils.py:27-28); a larger S keeps padding with the same values."""
    items = [getitem(int(i), dialogue_id, utterance_id, labels, text_emb, audio_emb) for i in indices]
    longest = max(t.shape[0] for t, _, _ in items)
    S = longest if S is None else int(S)
    assert S >= longest
    B = len(items)
    text = torch.zeros(B, S, items[0][0].shape[1], dtype=items[0][0].dtype)
    audio = torch.zeros(B, S, items[0][1].shape[1], dtype=items[0][1].dtype)
    emotion = torch.full((B, S), -1, dtype=torch.int64)
    for b, (t, a, e) in enumerate(items):
        text[b, :t.shape[0]] = t
        audio[b, :a.shape[0]] = a
        emotion[b, :e.shape[0]] = e
    return {"text": text, "audio": audio, "padding_mask": emotion == -1, "emotion": emotion}
