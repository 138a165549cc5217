"""GPU (-m gpu): the conditioning-latent encoder through the C ABI (tts_ar_load_conditioning_weights /
tts_ar_get_conditioning, UnifiedVoiceB200.get_conditioning) against (a) tests/golden/cond.npz -- outputs of the REFERENCE's
own UnifiedVoice.get_conditioning (oracle/make_golden_cond.py) -- and (b) the CPU oracle (oracle/cond_oracle.py) on other
shapes.  Tolerance: 2e-2 relative (max |diff| / max |ref|), the bf16 bar of BASELINE.json's north_star: the 1x1 convolutions of
the attention blocks run in bf16 with fp32 accumulation, everything else in fp32.
"""
import ast
import json
import os

import numpy as np
import pytest
import torch

from oracle.cond_oracle import OracleConditioning, make_cond_mel, make_cond_state_dict
from oracle.weights import ARConfig, make_inputs, make_state_dict

pytestmark = pytest.mark.gpu

BF16_REL = 2e-2
PARITY = {}


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module", autouse=True)
def _write_parity_record():
    yield
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out_dir = os.path.join(root, "gpurun_out") if os.path.isdir(os.path.join(root, "gpurun_out")) else os.path.join(root, "profiles")
    with open(os.path.join(out_dir, "parity_r02_cond.json"), "w") as f:
        json.dump(PARITY, f, indent=1, sort_keys=True)


def build_model(cond_seed, **kw):
    import __graft_entry__ as g
    g.build()
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    cfg = ARConfig(layers=1)
    sd = dict(make_state_dict(cfg, seed=1))
    sd.update(make_cond_state_dict(seed=cond_seed))
    return UnifiedVoiceB200(sd, ARGeometry(layers=1), max_utts=1, max_seqs=2, **kw), cfg, sd


@pytest.mark.parametrize("case", ["one_clip", "two_clips"])
def test_conditioning_latent_matches_the_reference_goldens(golden_dir, case):
    g = np.load(os.path.join(golden_dir, "cond.npz"))
    spec = ast.literal_eval(str(g[f"{case}_spec"]))
    m, cfg, sd = build_model(spec["w_seed"], max_text_len=32, max_new_tokens=200)
    mel = make_cond_mel(spec["batch"], spec["clips"], spec["frames"], seed=spec["in_seed"])
    lat = m.get_conditioning(mel.cuda())
    want = torch.from_numpy(g[f"{case}_latent"])
    e = rel(lat, want)
    PARITY[f"golden/{case}"] = dict(latent_rel=e, latent_abs_max=want.abs().max().item())
    print(f"conditioning {case}: rel vs reference {e:.3e}")
    assert lat.shape == want.shape and torch.isfinite(lat).all()
    assert e < BF16_REL
    if spec["clips"] == 1:                                   # [b, 80, T] input form (unified_voice.py:536)
        assert torch.equal(m.get_conditioning(mel[:, 0].cuda()), lat)
    m.close()


def test_conditioning_vs_oracle_shapes_and_use_as_prompt():
    """Frame counts around the tile edges (1, 63, 64, 65, 517 = a 6 s clip at hop 256) against the fp32 oracle, then the latent
    is used the way the reference uses it: as `speech_conditioning_latent` of inference_speech (a prompt pass must follow a
    get_conditioning call, which reuses the activation buffers)."""
    m, cfg, sd = build_model(41, max_text_len=16, max_new_tokens=600)
    o = OracleConditioning(sd)
    for frames, clips, seed in ((1, 1, 1), (63, 1, 2), (64, 2, 3), (65, 1, 4), (517, 2, 5)):
        mel = make_cond_mel(1, clips, frames, seed=seed)
        lat = m.get_conditioning(mel.cuda())
        want = o.get_conditioning(mel)
        e = rel(lat, want)
        PARITY[f"oracle/frames{frames}_clips{clips}"] = dict(latent_rel=e)
        print(f"conditioning frames={frames} clips={clips}: rel vs oracle {e:.3e}")
        assert e < BF16_REL
    with pytest.raises(ValueError):                          # capacity: more frames than activation rows
        m.get_conditioning(make_cond_mel(1, 1, 100000, seed=6).cuda())
    _, texts = make_inputs(cfg, 1, [16], seed=9)
    text = torch.stack(texts).cuda()
    codes = m.inference_speech(lat, text, num_return_sequences=2, max_generate_length=12, do_sample=False, ignore_eos=True)
    codes2 = m.inference_speech(lat, text, num_return_sequences=2, max_generate_length=12, do_sample=False, ignore_eos=True)
    assert codes.shape == (2, 12) and torch.equal(codes, codes2)
    m.close()


@pytest.mark.parametrize("case", ["short", "clip"])
def test_mel_spectrogram_matches_the_reference_goldens(golden_dir, case):
    """tts_ar_mel_spectrogram against the reference's TorchMelSpectrogram outputs (tests/golden/mel.npz): fp32 path (direct DFT
    with fp64 accumulation), so the bar is the fp32 one of BASELINE.json (1e-4 relative); then wav -> mel -> latent end to end
    against the two oracles chained."""
    from oracle.mel_oracle import make_wav, oracle_mel
    g = np.load(os.path.join(golden_dir, "mel.npz"))
    spec = ast.literal_eval(str(g[f"{case}_spec"]))
    norms = torch.from_numpy(g["mel_norms"])
    wav = make_wav(spec["batch"], spec["samples"], seed=spec["seed"])
    m, cfg, sd = build_model(43, max_text_len=16, max_new_tokens=600)
    mel = m.mel_spectrogram(wav.unsqueeze(1).cuda(), norms)
    want = torch.from_numpy(g[f"{case}_mel"])
    got = mel.cpu()[:, :, ::8] if case == "clip" else mel.cpu()
    e = rel(got, want)
    PARITY[f"mel/{case}"] = dict(mel_rel=e)
    print(f"mel {case}: rel vs reference {e:.3e}")
    assert got.shape == want.shape and e < 1e-4
    assert rel(m.mel_spectrogram(wav.cuda()), oracle_mel(wav)) < 1e-4          # no norms, [b, n] input form
    lat = m.get_conditioning(mel[:, None])                                       # emb/mel.py:92-97: [b, 1, 80, frames]
    want_lat = OracleConditioning(sd).get_conditioning(oracle_mel(wav, norms)[:, None])
    e2 = rel(lat, want_lat)
    PARITY[f"mel/{case}"]["latent_from_wav_rel"] = e2
    assert e2 < BF16_REL
    with pytest.raises(ValueError):
        m.mel_spectrogram(torch.zeros(1, 400).cuda())
    m.close()


def test_conditioning_bf16_source_weights():
    """TtsArCondWeights.dtype = 1 (bfloat16 state_dict): the other instantiation of the packing kernels."""
    import __graft_entry__ as g
    g.build()
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    cfg = ARConfig(layers=1)
    sd = dict(make_state_dict(cfg, seed=1))
    sd.update(make_cond_state_dict(seed=47))
    mel = make_cond_mel(1, 1, 90, seed=48)
    m32 = UnifiedVoiceB200(sd, ARGeometry(layers=1), max_utts=1, max_seqs=2, max_text_len=16, max_new_tokens=100)
    want = m32.get_conditioning(mel.cuda())
    m32.close()
    m16 = UnifiedVoiceB200({k: v.bfloat16() for k, v in sd.items()}, ARGeometry(layers=1), max_utts=1, max_seqs=2, max_text_len=16,
                           max_new_tokens=100)
    got = m16.get_conditioning(mel.cuda())
    m16.close()
    assert torch.isfinite(got).all() and rel(got, want) < BF16_REL
