"""CPU oracle of the mel-spectrogram front end of the conditioning path -- TEST INFRASTRUCTURE.

Restates TorchMelSpectrogram.forward (reference tortoise_tts/models/arch_utils.py:361-395, called by
tortoise_tts/emb/mel.py:50-65 format_autoregressive_conditioning): torchaudio.transforms.MelSpectrogram(n_fft=1024,
hop_length=256, win_length=1024, power=2, normalized=False, sample_rate=22050, f_min=0, f_max=8000, n_mels=80, norm="slaney")
-- a third-party dependency (torchaudio >= 0.13, setup.py:71; 2.11.0 in this image) whose published algorithm is:
  frames   centre-padded (reflect, n_fft // 2 each side), periodic Hann window, one-sided DFT of 1024 points, |.|^2
  filters  melscale_fbanks(513, 0, 8000, 80, 22050, norm="slaney", mel_scale="htk"): triangles between points equally spaced on
           the HTK mel scale m = 2595 log10(1 + f / 700), each scaled by 2 / (f[m+2] - f[m])
then log(clamp(mel, 1e-5)) / mel_norms[:, None] (arch_utils.py:389-394; mel_norms = data/models/mel_norms.pth, 80 values).
Pinned by tests/test_oracle_golden.py against tests/golden/mel.npz (oracle/make_golden_mel.py: the reference's own class).
"""
import math

import torch

N_FFT, HOP, N_MELS, SR, F_MIN, F_MAX = 1024, 256, 80, 22050, 0.0, 8000.0


def mel_filterbank():
    """[513, 80] fp32, torchaudio.functional.melscale_fbanks with norm='slaney', mel_scale='htk'."""
    n_freqs = N_FFT // 2 + 1
    all_freqs = torch.linspace(0, SR // 2, n_freqs)
    m_min = 2595.0 * math.log10(1.0 + F_MIN / 700.0)
    m_max = 2595.0 * math.log10(1.0 + F_MAX / 700.0)
    m_pts = torch.linspace(m_min, m_max, N_MELS + 2)
    f_pts = 700.0 * (10 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    down = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = torch.max(torch.zeros(1), torch.min(down, up))
    enorm = 2.0 / (f_pts[2: N_MELS + 2] - f_pts[:N_MELS])
    return fb * enorm.unsqueeze(0)


def make_wav(n_batch, n_samples, seed=0):
    """Speech-shaped synthetic audio in [-1, 1]: a few harmonics with slow amplitude modulation plus noise."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n_samples).float() / SR
    out = []
    for _ in range(n_batch):
        f0 = 90.0 + 120.0 * torch.rand(1, generator=g).item()
        w = sum((0.5 / (k + 1)) * torch.sin(2 * math.pi * f0 * (k + 1) * t + 6.28 * torch.rand(1, generator=g).item()) for k in range(8))
        env = 0.55 + 0.45 * torch.sin(2 * math.pi * 3.1 * t + 6.28 * torch.rand(1, generator=g).item())
        out.append(0.4 * w * env + 0.02 * torch.randn(n_samples, generator=g))
    return torch.stack(out).clamp(-1, 1)


def oracle_mel(wav, mel_norms=None):
    """wav [b, n] (or [b, 1, n]) -> [b, 80, 1 + n // 256]."""
    if wav.dim() == 3:
        wav = wav.squeeze(1)
    spec = torch.stft(wav.float(), N_FFT, hop_length=HOP, win_length=N_FFT, window=torch.hann_window(N_FFT), center=True,
                      pad_mode="reflect", normalized=False, onesided=True, return_complex=True)
    power = spec.abs().pow(2.0)                                        # [b, 513, frames]
    mel = torch.matmul(power.transpose(-1, -2), mel_filterbank()).transpose(-1, -2)
    mel = torch.log(torch.clamp(mel, min=1e-5))
    if mel_norms is not None:
        mel = mel / mel_norms.float().unsqueeze(0).unsqueeze(-1)
    return mel
