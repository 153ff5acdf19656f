"""The reference's src/train.py and src/test.py, UNCHANGED, run as __main__ (oracle/ref_harness.run_unchanged) on a tiny
synthetic MELD directory (SURVEY.md 8b "how the replacement is picked up", Appendix B).

CPU: (1) with the reference's own model the harness itself is sound (an epoch trains, a checkpoint is written, test.py loads
it); (2) with shim/model.py first on sys.path the drivers construct the B200 drop-in from config.yaml and -- there being no
GPU and no CPU path -- stop at the first forward with the product's RuntimeError, never falling back silently.
GPU: the same unchanged drivers train and evaluate the drop-in on the B200, and its checkpoint loads into the reference's
own module (strict) and back."""
import os

import pytest
import torch

from oracle import ref_harness as RH

needs_ref = pytest.mark.skipif(not RH.available(), reason="reference sources not present (run __graft_entry__.build() where /root/reference exists)")


@needs_ref
def test_unchanged_drivers_run_with_the_reference_model_on_cpu(tmp_path):
    scratch = RH.make_scratch(str(tmp_path / "meld"))
    r = RH.run_unchanged("train.py", scratch, b200_model=False, env={"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Training complete" in r.stdout and "Epoch: 0" in r.stdout
    assert os.path.exists(os.path.join(scratch, "checkpoints", "m2fnet.pth"))
    r = RH.run_unchanged("test.py", scratch, b200_model=False, env={"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Weighted_F1" in r.stdout


@needs_ref
def test_shim_puts_the_dropin_behind_the_unchanged_train_py(tmp_path):
    """No GPU visible: train.py falls back to device 'cpu' (src/train.py:20), builds the drop-in through `from model import
    M2FNet`, and the first forward raises the product's error -- there is no CPU path to fall back to."""
    scratch = RH.make_scratch(str(tmp_path / "meld"))
    r = RH.run_unchanged("train.py", scratch, b200_model=True, env={"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode != 0
    assert "Loaded 12 dialogues for training" in r.stdout          # the reference's Dataset ran (src/dataset.py:28)
    assert "no CPU path" in r.stderr and "M2FNet (B200-native)" in r.stderr


@pytest.mark.gpu
@needs_ref
def test_unchanged_drivers_train_and_test_the_dropin_on_the_gpu(tmp_path, pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    scratch = RH.make_scratch(str(tmp_path / "meld"), n_dialogues=(24, 8, 8), epochs=2)
    r = RH.run_unchanged("train.py", scratch, b200_model=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "Using device cuda:0" in r.stdout and "Training complete" in r.stdout and "Epoch: 1" in r.stdout
    ckpt = os.path.join(scratch, "checkpoints", "m2fnet.pth")
    assert os.path.exists(ckpt)
    r = RH.run_unchanged("test.py", scratch, b200_model=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "Accuracy=[" in r.stdout and "Testing complete" in r.stdout
    # the checkpoint the drop-in wrote is a plain reference checkpoint: strict load into the reference's own module
    _, ref_model = RH.load_reference()
    import yaml
    cfg = yaml.safe_load(open(os.path.join(scratch, "src", "config.yaml")))["model"]
    import sys
    sys.path.insert(0, RH.SHIM)
    from munch import Munch
    sys.path.remove(RH.SHIM)
    ref = ref_model.M2FNet(Munch.fromDict(cfg))
    sd = torch.load(ckpt, map_location="cpu")
    assert set(sd) == {"epoch", "model_state_dict", "optimizer_state_dict"}
    missing = ref.load_state_dict(sd["model_state_dict"], strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    # and the reference's CPU forward on that state_dict agrees with the drop-in's GPU forward (eval, fp32-parity mode)
    mine = pkg.M2FNet(Munch.fromDict(cfg), planes=3)
    mine.load_state_dict(sd["model_state_dict"], strict=True)
    mine = mine.to("cuda").eval()
    ref.eval()
    g = torch.Generator().manual_seed(0)
    text, audio = torch.randn(3, 5, 32, generator=g), torch.randn(3, 5, 16, generator=g)
    mask = torch.zeros(3, 5, dtype=torch.bool)
    mask[1, 3:] = True
    with torch.inference_mode():
        a = ref(text, audio, mask)
        b = mine(text.cuda(), audio.cuda(), mask.cuda()).cpu()
    assert torch.allclose(a[~mask], b[~mask], rtol=1e-4, atol=1e-5)
