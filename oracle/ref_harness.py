"""The UNMODIFIED reference, made runnable here and on the GPU box -- TEST / BASELINE INFRASTRUCTURE, not product code.

`build_ref()` (called by __graft_entry__.build()) copies the five files of the reference's fusion-model path
(src/{model,train,test,dataset,utils}.py + src/config.yaml) from /root/reference, byte for byte, into the git-ignored
`oracle/_ref/` (the GPU box has no /root/reference; `oracle/_ref/` travels with the snapshot like a built .so).  Nothing of
the reference is committed to this repository.

What uses it:
  * `bench.py --impl reference` and the in-line `cpu_baseline` leg: `load_reference()` imports the reference's own `train`
    module (its `train()` function, src/train.py:217-243) and `model` module (`M2FNet`, src/model.py) and times them on the
    host cores -- `cpu_baseline.kind = "reference"`;  also the stock-PyTorch-eager-on-the-B200 context numbers;
  * tests: `make_scratch()` writes a tiny MELD-shaped data directory (CSVs, embedding pickles, config.yaml) and
    `run_unchanged()` runs the reference's src/train.py / src/test.py as `__main__` in a fresh interpreter, with either the
    reference's own model or -- through shim/model.py -- the B200 drop-in on the import path.
"""
from __future__ import annotations

import importlib
import os
import pickle
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_SRC = "/root/reference/src"
REF_DIR = os.path.join(HERE, "_ref")
SHIM = os.path.join(ROOT, "shim")
FILES = ("model.py", "train.py", "test.py", "dataset.py", "utils.py", "config.yaml")


def build_ref(verbose: bool = False) -> bool:
    """Copy the reference's fusion-model sources into oracle/_ref/ (git-ignored).  Returns True when oracle/_ref is usable."""
    if os.path.isdir(REF_SRC):
        os.makedirs(REF_DIR, exist_ok=True)
        for f in FILES:
            shutil.copyfile(os.path.join(REF_SRC, f), os.path.join(REF_DIR, f))
        if verbose:
            print(f"oracle/_ref: copied {len(FILES)} files from {REF_SRC}")
    return available()


def ref_dir():
    """Directory holding the reference's src files: oracle/_ref (travels to the GPU box) or the reference checkout itself."""
    if all(os.path.exists(os.path.join(REF_DIR, f)) for f in FILES):
        return REF_DIR
    if all(os.path.exists(os.path.join(REF_SRC, f)) for f in FILES):
        return REF_SRC
    return None


def available() -> bool:
    return ref_dir() is not None


def load_reference():
    """Import the reference's `model` and `train` modules (unmodified files) in this process.  Returns (train_module or None,
    model_module).  `train` needs wandb / sklearn / tqdm / pandas (present in this image) and `munch` (shim/munch.py)."""
    d = ref_dir()
    if d is None:
        raise RuntimeError("the reference sources are not available (no oracle/_ref and no /root/reference)")
    import warnings
    warnings.filterwarnings("ignore")
    os.environ.setdefault("WANDB_MODE", "disabled")
    os.environ.setdefault("WANDB_SILENT", "true")
    added = [p for p in (d, SHIM) if p not in sys.path]
    for p in reversed(added):
        sys.path.insert(0, p)
    try:
        for name in ("model", "utils", "dataset", "train"):  # a stale same-named module must not shadow the reference's
            m = sys.modules.get(name)
            if m is not None and os.path.dirname(getattr(m, "__file__", "") or "") != d:
                del sys.modules[name]
        # shim/model.py must not win here: the reference's own model is wanted
        spec = importlib.util.spec_from_file_location("model", os.path.join(d, "model.py"))
        model_mod = importlib.util.module_from_spec(spec)
        sys.modules["model"] = model_mod
        spec.loader.exec_module(model_mod)
        try:
            train_mod = importlib.import_module("train")
        except Exception as e:  # noqa: BLE001  (e.g. wandb missing): the module alone is still the reference
            sys.stderr.write(f"ref_harness: could not import the reference's train.py ({e!r}); using its model with a restated loop\n")
            train_mod = None
    finally:
        for p in added:
            if p in sys.path:
                sys.path.remove(p)
    return train_mod, model_mod


def reference_config(**model_over):
    """`config.model` of the reference's src/config.yaml as the attribute dict its M2FNet expects."""
    import yaml
    sys.path.insert(0, SHIM)
    try:
        from munch import Munch
    finally:
        sys.path.remove(SHIM)
    d = ref_dir()
    with open(os.path.join(d, "config.yaml"), "rt", encoding="utf-8") as f:
        cfg = yaml.safe_load(f.read())["model"]
    for k, v in model_over.items():
        if isinstance(v, dict):
            cfg[k].update(v)
        else:
            cfg[k] = v
    return Munch.fromDict(cfg)


# ------------------------------------------------------------------------------------------ synthetic MELD directory
EMOTIONS = ("neutral", "joy", "sadness", "anger", "surprise", "fear", "disgust")


def make_scratch(path: str, n_dialogues=(12, 6, 6), max_len: int = 7, Dt: int = 32, Da: int = 16, seed: int = 0, model_over=None,
                 epochs: int = 1, batch_size: int = 4):
    """A directory the unchanged drivers run in: src/config.yaml (the reference's, with a small model / 1 epoch / no workers),
    data/MELD.Raw/{train,dev,test}_sent_emo.csv and embeddings/<kind>/{train,val,test}.pkl (one fp32 row per CSV row)."""
    import numpy as np
    import torch
    import yaml
    d = ref_dir()
    with open(os.path.join(d, "config.yaml"), "rt", encoding="utf-8") as f:
        cfg = yaml.safe_load(f.read())
    small = {"AUDIO": {"embedding_size": Da, "n_head": 2, "n_encoder_layers": 1}, "TEXT": {"embedding_size": Dt, "n_head": 2, "n_encoder_layers": 1},
             "FAM": {"embedding_size": 16, "n_head": 2, "n_layers": 1}, "CLASSIFIER": {"hidden_size": 16}}
    for k, v in {**small, **(model_over or {})}.items():
        if isinstance(v, dict):
            cfg["model"][k].update(v)
        else:
            cfg["model"][k] = v
    cfg["solver"]["epochs"] = epochs
    for split in ("train", "val", "test"):
        cfg[split]["data_loader"].update({"batch_size": batch_size, "num_workers": 0, "pin_memory": False})
    os.makedirs(os.path.join(path, "src"), exist_ok=True)
    with open(os.path.join(path, "src", "config.yaml"), "w") as f:
        yaml.safe_dump(cfg, f)
    rng = np.random.default_rng(seed)
    os.makedirs(os.path.join(path, "data", "MELD.Raw"), exist_ok=True)
    for split, csv, n in (("train", "train_sent_emo.csv", n_dialogues[0]), ("val", "dev_sent_emo.csv", n_dialogues[1]),
                          ("test", "test_sent_emo.csv", n_dialogues[2])):
        rows = []
        for did in range(n):
            for uid in range(int(rng.integers(1, max_len + 1))):
                rows.append((f"utterance {did}-{uid}", EMOTIONS[int(rng.integers(0, 7))], did, uid))
        with open(os.path.join(path, "data", "MELD.Raw", csv), "w") as f:
            f.write("Utterance,Emotion,Dialogue_ID,Utterance_ID\n")
            for r in rows:
                f.write(f"{r[0]},{r[1]},{r[2]},{r[3]}\n")
        g = torch.Generator().manual_seed(seed + len(rows))
        for kind, D, scale in ((cfg["embeddings"]["text"], Dt, 0.63), (cfg["embeddings"]["audio"], Da, 0.233)):
            os.makedirs(os.path.join(path, kind), exist_ok=True)
            with open(os.path.join(path, kind, f"{split}.pkl"), "wb") as f:
                pickle.dump(torch.randn(len(rows), D, generator=g) * scale, f)
    return path


def run_unchanged(script: str, scratch: str, b200_model: bool, timeout: int = 600, env=None):
    """Run the reference's src/<script> as __main__ in a fresh interpreter with cwd = scratch.  b200_model: shim/model.py
    precedes the reference's src on sys.path, so `from model import M2FNet` picks the drop-in; else the reference's own."""
    d = ref_dir()
    paths = [SHIM, d, ROOT] if b200_model else [d, SHIM, ROOT]  # `munch` always comes from shim/; `model` from whichever is first
    code = ("import sys, runpy, warnings\n"
            "warnings.filterwarnings('ignore')\n"
            f"sys.path[:0] = {paths!r}\n"
            f"runpy.run_path({os.path.join(d, script)!r}, run_name='__main__')\n")
    e = dict(os.environ, WANDB_MODE="disabled", WANDB_SILENT="true", **(env or {}))
    return subprocess.run([sys.executable, "-c", code], cwd=scratch, capture_output=True, text=True, timeout=timeout, env=e)


if __name__ == "__main__":
    print("oracle/_ref usable:", build_ref(verbose=True))
