"""Constant mel filterbank of the LogMel layer (espnet2/layers/log_mel.py:36-52).

The reference takes this table from ``librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=False)`` (Slaney mel scale,
Slaney area normalisation) and stores its transpose.  librosa is not a dependency here; this is the same published
construction in float64, rounded to fp32 once.  tests/test_abi.py checks it against the table frozen from the
reference run.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch


def _hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp, min_log_hz = 200.0 / 3, 1000.0
    logstep = math.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_hz / f_sp + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, f / f_sp)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp, min_log_hz = 200.0 / 3, 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_matrix(fs: int = 16000, n_fft: int = 512, n_mels: int = 80, fmin: float = 0.0, fmax: Optional[float] = None,
               device=None) -> torch.Tensor:
    """(n_fft//2 + 1, n_mels) fp32, contiguous (the layout b200slu_logmel_fwd reads)."""
    fmax = fs / 2 if fmax is None else fmax
    n_freq = n_fft // 2 + 1
    fft_f = np.linspace(0, fs / 2, n_freq)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fft_f[None, :]
    lower = -ramps[:-2] / fdiff[:-1, None]
    upper = ramps[2:] / fdiff[1:, None]
    w = np.maximum(0, np.minimum(lower, upper))
    w *= (2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels]))[:, None]
    t = torch.from_numpy(np.ascontiguousarray(w.T.astype(np.float32)))
    return t.to(device) if device is not None else t
