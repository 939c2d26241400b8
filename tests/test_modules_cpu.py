"""Host-side contract of the drop-in modules (no GPU): constructor keywords equal the reference's, state_dict names and
shapes equal the reference's (as restated by tests/test_step_gpu.make_sd, which the oracle consumes), unsupported
options fail loudly, and nothing runs on the CPU."""
import inspect
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import model as om
from test_step_gpu import make_sd

import espnet_slurp_b200.modules as m

# keyword lists of the reference constructors (conformer_encoder.py:82-111, transformer_decoder.py:233-250,
# ctc.py:19-28, espnet_model.py:42-69, default.py:25-41, specaug.py:27-40, utterance_mvn.py:24-30)
SIGS = {
    "ConformerEncoder": "input_size output_size attention_heads linear_units num_blocks dropout_rate positional_dropout_rate "
                        "attention_dropout_rate input_layer normalize_before concat_after positionwise_layer_type "
                        "positionwise_conv_kernel_size macaron_style rel_pos_type pos_enc_layer_type selfattention_layer_type "
                        "activation_type use_cnn_module zero_triu cnn_module_kernel padding_idx interctc_layer_idx "
                        "interctc_use_conditioning stochastic_depth_rate layer_drop_rate max_pos_emb_len",
    "TransformerDecoder": "vocab_size encoder_output_size attention_heads linear_units num_blocks dropout_rate "
                          "positional_dropout_rate self_attention_dropout_rate src_attention_dropout_rate input_layer "
                          "use_output_layer pos_enc_class normalize_before concat_after layer_drop_rate",
    "CTC": "odim encoder_output_size dropout_rate ctc_type reduce ignore_nan_grad zero_infinity",
    "ESPnetASRModel": "vocab_size token_list frontend specaug normalize preencoder encoder postencoder decoder ctc "
                      "joint_network ctc_weight interctc_weight ignore_id lsm_weight length_normalized_loss report_cer "
                      "report_wer sym_space sym_blank sym_sos sym_eos extract_feats_in_collect_stats lang_token_id",
    "DefaultFrontend": "fs n_fft win_length hop_length window center normalized onesided n_mels fmin fmax htk frontend_conf "
                       "apply_stft",
    "SpecAug": "apply_time_warp time_warp_window time_warp_mode apply_freq_mask freq_mask_width_range num_freq_mask "
               "apply_time_mask time_mask_width_range time_mask_width_ratio_range num_time_mask",
    "UtteranceMVN": "norm_means norm_vars eps",
    "GlobalMVN": "stats_file norm_means norm_vars eps",
}


@pytest.mark.parametrize("name", sorted(SIGS))
def test_constructor_keywords_match_reference(name):
    got = [p for p in inspect.signature(getattr(m, name).__init__).parameters if p != "self"]
    assert got == SIGS[name].split()
    ref_root = "/root/reference"
    if os.path.isdir(ref_root):  # cross-check the list itself against the reference source when it is around
        path = {"ConformerEncoder": "espnet2/asr/encoder/conformer_encoder.py", "TransformerDecoder":
                "espnet2/asr/decoder/transformer_decoder.py", "CTC": "espnet2/asr/ctc.py", "ESPnetASRModel":
                "espnet2/asr/espnet_model.py", "DefaultFrontend": "espnet2/asr/frontend/default.py", "SpecAug":
                "espnet2/asr/specaug/specaug.py", "UtteranceMVN": "espnet2/layers/utterance_mvn.py",
                "GlobalMVN": "espnet2/layers/global_mvn.py"}[name]
        import ast
        tree = ast.parse(open(os.path.join(ref_root, path)).read())
        cls = [n for n in ast.walk(tree) if isinstance(n, ast.ClassDef) and n.name == name][-1]
        init = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "__init__"][0]
        assert [a.arg for a in init.args.args if a.arg != "self"] == got


def test_state_dict_names_and_shapes():
    enc = m.ConformerEncoder(80, 64, 4, 128, 2, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=7)
    dec = m.TransformerDecoder(30, 64, 4, 128, 2)
    ctc = m.CTC(30, 64)
    tokens = ["<blank>"] + ["x"] * 28 + ["<sos/eos>"]
    model = m.ESPnetASRModel(30, tokens, m.DefaultFrontend(), m.SpecAug(time_mask_width_range=40), m.UtteranceMVN(), None, enc,
                             None, dec, ctc, None, ctc_weight=0.3, lsm_weight=0.1)
    want = make_sd(om.EncCfg(80, 64, 4, 128, 2, 7, "latest"), om.DecCfg(30, 64, 4, 128, 2), 30)
    got = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    assert got.pop("frontend.logmel.melmat") == (257, 80)  # log_mel.py:52
    params = {k: v for k, v in got.items() if not k.endswith(("running_mean", "running_var", "num_batches_tracked"))}
    assert params == {k: tuple(v.shape) for k, v in want.items()}
    bn = sorted(k for k in got if k not in params)
    assert bn[:3] == ["encoder.encoders.0.conv_module.norm.num_batches_tracked", "encoder.encoders.0.conv_module.norm.running_mean",
                      "encoder.encoders.0.conv_module.norm.running_var"]
    assert enc.output_size() == 64 and enc.interctc_use_conditioning is False
    # default init follows torch's layers (Linear kaiming-uniform, LayerNorm ones/zeros, pos_bias xavier-uniform)
    assert float(enc.after_norm.weight.min()) == 1.0 and float(enc.after_norm.bias.abs().max()) == 0.0
    assert enc.after_norm.eps == 1e-12
    u = enc.encoders[0].self_attn.pos_bias_u
    assert float(u.abs().max()) <= (6.0 / (4 + 16)) ** 0.5 + 1e-6 and float(u.abs().max()) > 0


def test_unsupported_options_raise():
    with pytest.raises(NotImplementedError):
        m.ConformerEncoder(80, 64, 4, 128, 2, macaron_style=False)
    with pytest.raises(NotImplementedError):
        m.ConformerEncoder(80, 64, 4, 128, 2, macaron_style=True, input_layer="linear")
    enc_sc = m.ConformerEncoder(80, 64, 4, 128, 2, macaron_style=True, interctc_layer_idx=[1], interctc_use_conditioning=True)
    assert enc_sc._cfg.interctc_use_conditioning and enc_sc.conditioning_layer is None  # the model creates it (:96-101)
    with pytest.raises(NotImplementedError):  # stochastic depth: the reference's own EncoderLayer takes no such argument
        m.ConformerEncoder(80, 64, 4, 128, 2, macaron_style=True, stochastic_depth_rate=0.1)
    with pytest.raises(AssertionError):  # conformer_encoder.py:284-285: taps must lie strictly inside the stack
        m.ConformerEncoder(80, 64, 4, 128, 2, macaron_style=True, interctc_layer_idx=[2])
    assert m.ConformerEncoder(80, 64, 4, 128, 3, macaron_style=True, interctc_layer_idx=[2, 1])._cfg.interctc_layer_idx == (1, 2)
    with pytest.raises(NotImplementedError):
        m.TransformerDecoder(30, 64, concat_after=True)
    with pytest.raises(NotImplementedError):
        m.CTC(30, 64, ctc_type="gtnctc")
    with pytest.raises(ValueError):
        m.SpecAug(apply_time_warp=False, apply_freq_mask=False, apply_time_mask=False)
    with pytest.raises(NotImplementedError):
        m.UtteranceMVN(norm_vars=True)


def test_no_cpu_fallback():
    enc = m.ConformerEncoder(80, 64, 4, 128, 1, macaron_style=True, rel_pos_type="latest")
    with pytest.raises(RuntimeError, match="CUDA"):
        enc(torch.randn(2, 50, 80), torch.tensor([50, 40]))
    with pytest.raises(RuntimeError, match="CUDA"):
        m.DefaultFrontend()(torch.randn(2, 4000), torch.tensor([4000, 3000]))
    with pytest.raises(RuntimeError, match="CUDA"):
        m.CTC(10, 64)(torch.randn(2, 20, 64), torch.tensor([20, 10]), torch.ones(2, 3).long(), torch.tensor([3, 2]))


def test_specaug_draws_follow_reference_ranges():
    g = torch.Generator().manual_seed(0)
    flens = torch.tensor([300, 40, 9])
    d = m.specaug_draws(flens, 300, 80, 5, (0, 30), 2, (0, 40), 2, gen=g)
    c, w = d["center"].tolist(), d["warped"].tolist()
    assert 5 <= c[0] < 295 and c[0] - 5 < w[0] <= c[0] + 5       # time_warp.py:28-29
    assert c[2] == -1                                              # too short to warp: passes through (time_warp.py:26)
    f, t = d["fmask"], d["tmask"]
    assert f.shape == (3, 2, 2) and int(f[..., 1].max()) < 30 and int((f[..., 0] + f[..., 1]).max()) <= 80 + 30
    assert t.shape == (3, 2, 2) and int(t[..., 1].max()) < 40


def test_reference_checkpoint_loads_strictly():
    """A state_dict saved by the unmodified reference model (tests/golden/decode.npz, frontend + BatchNorm buffers
    included) loads with strict=True, and the filterbank buffer equals the reference's librosa table."""
    import numpy as np
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "decode.npz"))
    sd = {k[3:]: torch.tensor(z[k]) for k in z.files if k.startswith("sd/")}
    V = 20
    enc = m.ConformerEncoder(80, 16, 2, 32, 2, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=7)
    dec = m.TransformerDecoder(V, 16, 2, 32, 2)
    fe = m.DefaultFrontend()
    assert torch.equal(fe.logmel.melmat, sd["frontend.logmel.melmat"])
    model = m.ESPnetASRModel(V, ["<blank>", "<unk>"] + [f"t{i}" for i in range(V - 3)] + ["<sos/eos>"], fe,
                             m.SpecAug(time_mask_width_range=(0, 40)), m.UtteranceMVN(), None, enc, None, dec, m.CTC(V, 16), None)
    res = model.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert sorted(model.state_dict()) == sorted(sd)
    import copy
    import pickle
    clone = copy.deepcopy(model)  # the owner links / flat buffers are rebuilt lazily, never copied or pickled
    assert "_owner" not in clone.decoder.__dict__ and model.decoder.__dict__["_owner"]() is model
    pickle.loads(pickle.dumps(model))
