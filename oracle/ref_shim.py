"""Import the UNMODIFIED reference (/root/reference) in the build container.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Used only by oracle/make_golden.py
and by tests that are skipped when /root/reference is absent (it does not exist on
the GPU box).  No reference file is edited or copied: missing third-party modules
are stubbed and three name bindings are patched in the importing modules'
namespaces, as recorded in SURVEY.md section 8(c).
"""
from __future__ import annotations

import os
import sys
import types

REF_ROOT = os.environ.get("ESPNET_REF_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "espnet2"))


_installed = False


def install():
    """Make ``import espnet2...`` work against REF_ROOT.  Idempotent."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError(f"reference not found at {REF_ROOT}")
    import numpy as np
    import torch

    # (1) typeguard 4.x rejects the reference's `x: int = None` annotations.
    import typeguard

    typeguard.check_argument_types = lambda *a, **k: True
    typeguard.check_return_type = lambda *a, **k: True

    # (2) stubs for absent third-party modules (import-only on this path).
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules.setdefault(name, m)
        return sys.modules[name]

    def parse_size(s):
        if isinstance(s, (int, float)):
            return int(s)
        s = str(s).strip().lower()
        mult = 1
        if s.endswith("k"):
            mult, s = 1000, s[:-1]
        return int(float(s) * mult)

    stub("humanfriendly", parse_size=parse_size, format_timespan=lambda x: str(x))

    class ComplexTensor:  # holder with .real/.imag/.dim(), enough for default.py
        def __init__(self, real, imag):
            self.real, self.imag = real, imag

        def dim(self):
            return self.real.dim()

        def size(self, *a):
            return self.real.size(*a)

        @property
        def shape(self):
            return self.real.shape

    tc = stub("torch_complex")
    tct = stub("torch_complex.tensor", ComplexTensor=ComplexTensor)
    tc.tensor = tct
    tc.ComplexTensor = ComplexTensor
    stub("torch_complex.functional")
    stub("editdistance", eval=lambda a, b: 0)
    stub("pytorch_wpe", wpe_one_iteration=None)

    from oracle.model import slaney_mel_matrix

    def mel(sr, n_fft, n_mels=80, fmin=0.0, fmax=None, htk=False, **kw):
        assert not htk
        return slaney_mel_matrix(sr, n_fft, n_mels, fmin, fmax).T.astype(np.float32)

    lib = stub("librosa")
    lib.filters = stub("librosa.filters", mel=mel)

    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)

    # (3) names the newer espnet2/ expects from the older espnet/.
    import espnet.nets.pytorch_backend.transformer.subsampling as sub

    if not hasattr(sub, "Conv2dSubsampling2"):
        class Conv2dSubsampling2(torch.nn.Module):  # never instantiated on this path
            pass

        sub.Conv2dSubsampling2 = Conv2dSubsampling2
    try:
        import espnet.nets.pytorch_backend.transducer.utils as tu

        for n in ("select_k_expansions", "subtract", "is_prefix", "recombine_hyps"):
            if not hasattr(tu, n):
                setattr(tu, n, None)
    except Exception:
        pass

    # (4) signature drift, patched in the IMPORTING modules' namespaces only.
    import espnet.nets.pytorch_backend.nets_utils as nu
    from espnet.nets.pytorch_backend.transformer.repeat import repeat as _repeat

    def repeat(N, fn, layer_drop_rate=0.0):
        return _repeat(N, fn)

    def make_pad_mask(lengths, xs=None, length_dim=-1, maxlen=None):
        m = nu.make_pad_mask(lengths, xs, length_dim)
        if maxlen is not None and xs is None and m.size(-1) < maxlen:
            m = torch.nn.functional.pad(m, (0, maxlen - m.size(-1)), value=True)
        return m

    import espnet2.asr.encoder.conformer_encoder as ce

    _EL = ce.EncoderLayer

    def EncoderLayer(*a, **k):
        if len(a) == 9:  # trailing stochastic_depth_rate
            assert a[-1] == 0.0
            a = a[:-1]
        k.pop("stochastic_depth_rate", None)
        return _EL(*a, **k)

    ce.EncoderLayer = EncoderLayer
    ce.repeat = repeat
    import espnet2.asr.decoder.transformer_decoder as td

    td.repeat = repeat
    td.make_pad_mask = make_pad_mask
    try:
        import espnet2.asr.encoder.transformer_encoder as te

        te.repeat = repeat
    except Exception:
        pass
    _installed = True


def build_reference_model(vocab: int, enc_conf: dict, dec_conf: dict, ctc_weight=0.3, lsm_weight=0.1,
                          frontend=True, specaug_conf=None, interctc_weight=0.0):
    """The reference's ESPnetASRModel, assembled as espnet2/tasks/asr.py:439-562 does."""
    install()
    from espnet2.asr.ctc import CTC
    from espnet2.asr.decoder.transformer_decoder import TransformerDecoder
    from espnet2.asr.encoder.conformer_encoder import ConformerEncoder
    from espnet2.asr.espnet_model import ESPnetASRModel
    from espnet2.asr.frontend.default import DefaultFrontend
    from espnet2.asr.specaug.specaug import SpecAug
    from espnet2.layers.utterance_mvn import UtteranceMVN

    fe = DefaultFrontend(frontend_conf=None) if frontend else None  # WPE/MVDR off by default: no-op
    input_size = 80
    enc = ConformerEncoder(input_size=input_size, **enc_conf)
    dec = TransformerDecoder(vocab_size=vocab, encoder_output_size=enc.output_size(), **dec_conf)
    ctc = CTC(odim=vocab, encoder_output_size=enc.output_size())
    token_list = ["<blank>", "<unk>"] + [f"t{i}" for i in range(vocab - 3)] + ["<sos/eos>"]
    model = ESPnetASRModel(
        vocab_size=vocab, token_list=token_list, frontend=fe,
        specaug=SpecAug(**specaug_conf) if specaug_conf else None,
        normalize=UtteranceMVN() if frontend else None, preencoder=None, encoder=enc, postencoder=None,
        decoder=dec, ctc=ctc, joint_network=None, ctc_weight=ctc_weight, interctc_weight=interctc_weight, lsm_weight=lsm_weight,
        length_normalized_loss=False, report_cer=False, report_wer=False,
    )
    return model
