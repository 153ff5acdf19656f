"""Host-side mirror of the reference interface (no GPU): constructor checks, state_dict names/shapes, shared final
norm aliases, strict checkpoint interchange with the reference module, CPU tensors refused."""
import importlib.util
import os
import warnings

import pytest
import torch

from oracle import m2fnet_oracle as O

REF = "/root/reference/src/model.py"


def test_state_dict_matches_reference_names_and_shapes(pkg):
    for over in (dict(), dict(AUDIO=dict(n_transformers=2, embedding_size=300, n_head=10), TEXT=dict(n_transformers=3, embedding_size=1024, n_head=16),
                              FAM=dict(embedding_size=1024, n_head=16, n_layers=2), CLASSIFIER=dict(hidden_size=1024, output_size=7, n_layers=3)),
                 dict(TEXT=dict(enabled=False), FAM=dict(enabled=False))):
        cfg = O.default_model_cfg(**over)
        m = pkg.M2FNet(cfg)
        sd, want = m.state_dict(), O.param_shapes(cfg)
        assert set(sd) == set(want)
        assert all(tuple(sd[k].shape) == want[k] for k in want)
        n_unique = len({O.canonical_name(k) for k in want})
        assert len(list(m.parameters())) == n_unique


def test_default_config_counts(pkg):
    m = pkg.M2FNet(O.default_model_cfg())
    assert len(m.state_dict()) == 186 and sum(p.numel() for p in m.parameters()) == 86_251_783


def test_shared_final_norm_is_one_tensor(pkg):
    m = pkg.M2FNet(O.default_model_cfg(AUDIO=dict(n_transformers=2), TEXT=dict(n_transformers=2)))
    sd = m.state_dict()
    assert sd["audio_encoders.0.norm.weight"].data_ptr() == sd["audio_encoders.1.norm.weight"].data_ptr()
    assert len(sd) == 334 and len(list(m.parameters())) == 330


def test_constructor_errors(pkg):
    with pytest.raises(ValueError, match="At least one of audio and text must be enabled!"):
        pkg.M2FNet(O.default_model_cfg(AUDIO=dict(enabled=False), TEXT=dict(enabled=False), FAM=dict(enabled=False)))
    with pytest.raises(ValueError, match="Fusion Attention Module can only be used with both audio and text enabled!"):
        pkg.M2FNet(O.default_model_cfg(AUDIO=dict(enabled=False)))
    with pytest.raises(AssertionError, match="embed_dim must be divisible by num_heads"):
        pkg.M2FNet(O.default_model_cfg(AUDIO=dict(embedding_size=300)))  # 300 % 8 != 0


def test_encoder_layers_start_as_clones(pkg):
    m = pkg.M2FNet(O.default_model_cfg())
    sd = m.state_dict()
    assert torch.equal(sd["text_encoders.0.layers.0.linear1.weight"], sd["text_encoders.0.layers.5.linear1.weight"])
    assert not torch.equal(sd["text_encoders.0.layers.0.linear1.weight"], sd["audio_encoders.0.layers.0.linear1.weight"])
    assert sd["text_encoders.0.layers.0.self_attn.in_proj_bias"].abs().max() == 0


def test_parameters_are_views_of_one_arena_and_survive_load(pkg):
    cfg = O.default_model_cfg(**{k: v for k, v in [("AUDIO", dict(embedding_size=32, n_head=4, n_encoder_layers=1)),
                                                   ("TEXT", dict(embedding_size=32, n_head=4, n_encoder_layers=1)),
                                                   ("FAM", dict(embedding_size=32, n_head=4, n_layers=1)),
                                                   ("CLASSIFIER", dict(hidden_size=16, output_size=7, n_layers=2))]})
    m = pkg.M2FNet(cfg)
    params = {k: v.float() for k, v in O.synth_params(cfg).items()}
    m.load_state_dict(params, strict=True)
    assert m._views_intact()
    for name, p in m.named_parameters():
        assert torch.equal(p, params[name])
    lo = min(o for _, o, _ in m._flat)
    assert lo == 0 and m.flat_parameters().numel() % 64 == 0


def test_cpu_forward_is_refused(pkg):
    m = pkg.M2FNet(O.default_model_cfg(TEXT=dict(enabled=False), FAM=dict(enabled=False)))
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.zeros(1, 2, 768), torch.zeros(1, 2, 768), torch.zeros(1, 2, dtype=torch.bool))


@pytest.mark.skipif(not os.path.exists(REF), reason="reference tree only exists in the build container")
def test_strict_checkpoint_interchange_with_reference(pkg):
    spec = importlib.util.spec_from_file_location("ref_model", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    cfg = O.default_model_cfg(AUDIO=dict(embedding_size=32, n_head=4, n_transformers=2, n_encoder_layers=1),
                              TEXT=dict(embedding_size=48, n_head=4, n_encoder_layers=2), FAM=dict(embedding_size=32, n_head=4, n_layers=2),
                              CLASSIFIER=dict(hidden_size=16, output_size=7, n_layers=3))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = ref.M2FNet(cfg)
    ours = pkg.M2FNet(cfg)
    assert not ours.load_state_dict(r.state_dict(), strict=True).missing_keys
    assert not r.load_state_dict(ours.state_dict(), strict=True).missing_keys
    assert sorted(n for n, _ in r.named_parameters()) == sorted(n for n, _ in ours.named_parameters())


def test_state_dict_keys_survive_a_forward_setup(pkg):
    """Internal caches must not leak into state_dict (nn.Module registers any Parameter assigned as an attribute)."""
    import importlib
    D = importlib.import_module(pkg.__name__ + ".data")
    m = pkg.M2FNet(D.default_model_config(AUDIO=dict(embedding_size=32, n_head=2, n_encoder_layers=1),
                                          TEXT=dict(embedding_size=32, n_head=2, n_encoder_layers=1),
                                          FAM=dict(embedding_size=32, n_head=2, n_layers=1),
                                          CLASSIFIER=dict(hidden_size=32, output_size=7, n_layers=2)))
    keys = set(m.state_dict().keys())
    assert m._grad_anchor() is not None
    assert set(m.state_dict().keys()) == keys
    assert all(not k.startswith("_") for k in keys)


def test_inference_workspace_is_recycled(pkg):
    """Plan geometry is host arithmetic (no device needed): an inference plan shares per-layer buffers, alternates two output slots
    and lays the fusion / classifier buffers over the finished audio stack (csrc/plan.cu, m2f_plan_create), so the C5 batch
    (512 dialogues x 256 utterances, /root/reference/src/test.py:51-74 at BASELINE.json configs[4]) needs < 8 GB in bf16 mode --
    and far less than the training plan of the same geometry, which keeps every activation for the backward."""
    import ctypes as C
    L = pkg._lib.lib()
    sizes = {}
    for prec in ("bf16", "fp32"):
        m = pkg.M2FNet(O.default_model_cfg(dropout=0.4), precision=prec)
        for B, S, training in ((512, 256, 0), (64, 33, 0), (64, 33, 1)):
            h = C.c_void_p()
            assert L.m2f_plan_create(m._handle, B, S, training, C.byref(h)) == 0
            sizes[prec, B, S, training] = int(L.m2f_plan_workspace_bytes(h))
            L.m2f_plan_destroy(h)
    assert sizes["bf16", 512, 256, 0] < 8e9
    assert sizes["fp32", 512, 256, 0] < 16e9
    assert sizes["bf16", 64, 33, 0] * 6 < sizes["bf16", 64, 33, 1]
