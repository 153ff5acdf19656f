"""Writes tests/golden/collate_ref.npz by running the reference's OWN Dataset + collate_fn (src/dataset.py) on a small
synthetic MELD-shaped split.  Runs in the build container only (needs /root/reference); the fixture travels.

    python oracle/gen_golden_collate.py

The scratch tree the reference reads (./src/config.yaml, ./data/MELD.Raw/*.csv, ./embeddings/*/val.pkl) and a ~10-line
stand-in for the `munch` package (absent from the image; the reference only uses Munch.fromYAML + attribute access)
are created under oracle/_scratch_collate/ and removed afterwards.
"""
import os
import pickle
import shutil
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference/src"
OUT = os.path.join(ROOT, "tests", "golden", "collate_ref.npz")
EMOTIONS = ["neutral", "joy", "sadness", "anger", "surprise", "fear", "disgust"]  # src/dataset.py:22

MUNCH_SHIM = '''
import yaml
class Munch(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)
    @classmethod
    def wrap(cls, d):
        if isinstance(d, dict):
            return cls({k: cls.wrap(v) for k, v in d.items()})
        return d
    @classmethod
    def fromYAML(cls, s):
        return cls.wrap(yaml.safe_load(s))
'''


def main():
    import pandas as pd
    scratch = os.path.join(ROOT, "oracle", "_scratch_collate")
    shutil.rmtree(scratch, ignore_errors=True)
    for d in ("src", "data/MELD.Raw", "embeddings/text", "embeddings/audio", "shim"):
        os.makedirs(os.path.join(scratch, d))
    rng = np.random.default_rng(7)
    # dialogues appear interleaved and with their utterances out of order; ids are not contiguous; the val split drops
    # the row (Dialogue_ID 110, Utterance_ID 7) before the embeddings are indexed (src/utils.py:54-56)
    lengths = {4: 5, 110: 9, 2: 1, 57: 12, 9: 3, 300: 7, 12: 2}
    rows = [(d, u) for d, n in lengths.items() for u in range(n)]
    order = rng.permutation(len(rows))
    rows = [rows[i] for i in order]
    df = pd.DataFrame({"Sr No.": np.arange(len(rows)), "Utterance": [f"u{d}_{u}" for d, u in rows],
                       "Emotion": [EMOTIONS[int(rng.integers(0, 7))] for _ in rows],
                       "Dialogue_ID": [d for d, _ in rows], "Utterance_ID": [u for _, u in rows]})
    df.to_csv(os.path.join(scratch, "data/MELD.Raw/dev_sent_emo.csv"), index=False)
    n_kept = len(rows) - 1
    Dt, Da = 8, 6
    text_emb = torch.from_numpy(rng.standard_normal((n_kept, Dt)).astype(np.float32))
    audio_emb = torch.from_numpy(rng.standard_normal((n_kept, Da)).astype(np.float32))
    pickle.dump(text_emb, open(os.path.join(scratch, "embeddings/text/val.pkl"), "wb"))
    pickle.dump(audio_emb, open(os.path.join(scratch, "embeddings/audio/val.pkl"), "wb"))
    open(os.path.join(scratch, "src/config.yaml"), "w").write("embeddings:\n  text: embeddings/text\n  audio: embeddings/audio\n")
    open(os.path.join(scratch, "shim/munch.py"), "w").write(MUNCH_SHIM)
    cwd = os.getcwd()
    os.chdir(scratch)
    sys.path.insert(0, REF_SRC)
    sys.path.insert(0, os.path.join(scratch, "shim"))
    try:
        import dataset as ref_dataset  # the reference's module, unmodified
        ds = ref_dataset.Dataset("val")
        frame = ds.text  # after the corrupted-row filter + reset_index: row i <-> embedding row i
        blob = {"dialogue_id": frame["Dialogue_ID"].to_numpy().astype(np.int64), "utterance_id": frame["Utterance_ID"].to_numpy().astype(np.int64),
                "labels": ds.get_labels().astype(np.int64), "text_emb": text_emb.numpy(), "audio_emb": audio_emb.numpy(),
                "dataset_dialogue_ids": np.asarray(ds.dialogue_ids).astype(np.int64)}
        batches = [[0, 1, 2, 3], [4, 5, 6], [2], [6, 0, 3, 3, 1]]
        for bi, idx in enumerate(batches):
            out = ref_dataset.collate_fn([ds[i] for i in idx])
            blob[f"batch{bi}/indices"] = np.asarray(idx, dtype=np.int64)
            for k, v in out.items():
                blob[f"batch{bi}/{k}"] = v.numpy()
            print(f"batch {bi}: indices {idx} -> text {tuple(out['text'].shape)}, valid {int((~out['padding_mask']).sum())}")
        blob["n_batches"] = np.array(len(batches))
    finally:
        os.chdir(cwd)
        shutil.rmtree(scratch, ignore_errors=True)
    np.savez_compressed(OUT, **blob)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
