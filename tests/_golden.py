"""Helpers shared by the golden-fixture tests (CPU oracle pin and GPU model parity)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_case(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def grad_names(z):
    return sorted({k.split("/")[1] for k in z if k.startswith("grad/")})


def grad_errors(z, grads):
    """Per-tensor relative L2 error of `grads` (dict canonical name -> tensor) against the fixture, plus the error
    of the concatenation of everything the fixture stores (full tensors, or strided samples of the large ones)."""
    per, num, den = {}, 0.0, 0.0
    for n in grad_names(z):
        g = grads[n].detach().double().flatten().cpu()
        if f"grad/{n}/full" in z:
            ref = torch.from_numpy(z[f"grad/{n}/full"])
            got = g
        else:
            step = int(z[f"grad/{n}/step"])
            ref = torch.from_numpy(z[f"grad/{n}/sample"])
            got = g[::step]
        d = torch.linalg.vector_norm(got - ref).item()
        r = torch.linalg.vector_norm(ref).item()
        per[n] = d / r if r > 0 else d
        num += d * d
        den += r * r
    return per, (num / den) ** 0.5 if den > 0 else num ** 0.5


def case_inputs(z, cfg, B, S, ragged):
    """(text, audio) fp64 tensors of a case: stored in the fixture, or regenerated when the case is RNG-free."""
    if ragged == "rngfree":
        from oracle import m2fnet_oracle as O
        b = O.synth_inputs(B, S, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size, C=cfg.CLASSIFIER.output_size)
        assert torch.equal(b["padding_mask"], torch.from_numpy(z["mask"])) and torch.equal(b["emotion"], torch.from_numpy(z["emotion"]))
        return b["text"].double(), b["audio"].double()
    return torch.from_numpy(z["text"]), torch.from_numpy(z["audio"])
