"""Imports the reference's AR classes from /root/reference BY PATH (TEST INFRASTRUCTURE; build
container only -- /root/reference does not exist on the GPU box, nothing run there imports this).

`import tortoise_tts` fails here (h5py / soundfile / librosa missing) and the reference's
generate() override does not run on transformers 5.5.0, so this shim
(a) registers empty package stubs so `tortoise_tts/__init__` chains are not executed,
(b) stubs librosa and the HF symbols the reference imports but this path never calls,
(c) exposes `ref_generate()`: the reference's own UnifiedVoice.compute_embeddings +
    GPT2InferenceModel.{prepare_inputs_for_generation, forward} driven by a loop that follows the
    reference's sample_stream (stream_generator.py:1106-1190) with HF's own processor classes.
"""
import sys
import types
import importlib
import importlib.machinery

REF = "/root/reference/tortoise_tts"
_uv = None


def load_reference():
    global _uv
    if _uv is not None:
        return _uv
    sys.dont_write_bytecode = True

    def stub(name):
        m = types.ModuleType(name)
        m.__spec__ = importlib.machinery.ModuleSpec(name, None)
        sys.modules[name] = m
        return m

    import transformers
    from transformers import GPT2Config, GPT2Model, GPT2PreTrainedModel, LogitsProcessorList  # noqa
    from transformers import GenerationMixin, GenerationConfig, StoppingCriteriaList, PreTrainedModel  # noqa
    import transformers.generation.utils as gu
    import transformers.generation.logits_process  # noqa
    import transformers.modeling_outputs  # noqa
    lib, util, filt = stub("librosa"), stub("librosa.util"), stub("librosa.filters")
    lib.util, lib.filters = util, filt
    util.pad_center = util.tiny = util.normalize = filt.mel = lambda *a, **k: None
    T = sys.modules["transformers"]
    for n in ["DisjunctiveConstraint", "BeamSearchScorer", "PhrasalConstraint", "ConstrainedBeamSearchScorer"]:
        if n not in T.__dict__:
            setattr(T, n, type(n, (), {}))
    if not hasattr(gu, "SampleOutput"):
        gu.SampleOutput = gu.GenerateOutput
    mp = stub("transformers.utils.model_parallel_utils")
    mp.get_device_map = None
    mp.assert_device_map = None
    for pkg, path in [("tortoise_tts", REF), ("tortoise_tts.models", REF + "/models")]:
        m = types.ModuleType(pkg)
        m.__path__ = [path]
        sys.modules[pkg] = m
    _uv = importlib.import_module("tortoise_tts.models.unified_voice")
    return _uv


def build_reference_model(sd, cfg):
    """Instantiate the reference UnifiedVoice at cfg's size and load the synthetic weights
    (strict=False like the reference's own loader, models/__init__.py:105: the conditioning
    encoder / text_head are not on this path and keep their default init)."""
    import torch
    uv = load_reference()
    m = uv.UnifiedVoice(layers=cfg.layers, model_dim=cfg.model_dim, heads=cfg.heads,
                        max_text_tokens=cfg.max_text_tokens, max_mel_tokens=cfg.max_mel_tokens,
                        max_conditioning_inputs=cfg.max_conditioning_inputs,
                        number_text_tokens=cfg.number_text_tokens, start_text_token=cfg.start_text_token,
                        number_mel_codes=cfg.number_mel_codes, start_mel_token=cfg.start_mel_token,
                        stop_mel_token=cfg.stop_mel_token, checkpointing=False)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.startswith(("conditioning_encoder.", "text_head.")) for k in missing), missing
    m.eval()
    return m


def ref_generate(m, latent, text, n_cand=1, max_new=8, do_sample=False, temperature=1.0, top_k=0,
                 top_p=1.0, repetition_penalty=1.0, ignore_eos=False, want_logits=0, seed=0):
    """The reference's classes, stepped as sample_stream does.  Returns dict like OracleAR.generate."""
    import torch
    from transformers.generation.logits_process import (
        LogitsProcessorList, RepetitionPenaltyLogitsProcessor, TemperatureLogitsWarper,
        TopKLogitsWarper, TopPLogitsWarper)
    with torch.inference_mode():
        ids = m.compute_embeddings(latent, text)                  # [U, P+1]
        im = m.inference_model
        ids = ids.repeat_interleave(n_cand, 0)                    # _expand_inputs_for_generation
        P1 = ids.shape[1]
        procs = LogitsProcessorList()
        if repetition_penalty != 1.0:
            procs.append(RepetitionPenaltyLogitsProcessor(penalty=repetition_penalty))
        warp = LogitsProcessorList()                              # ref _get_logits_warper order
        if do_sample:
            if temperature != 1.0:
                warp.append(TemperatureLogitsWarper(temperature))
            if top_k:
                warp.append(TopKLogitsWarper(top_k=top_k, min_tokens_to_keep=1))
            if top_p < 1.0:
                warp.append(TopPLogitsWarper(top_p=top_p, min_tokens_to_keep=1))
        torch.manual_seed(seed)
        eos = pad = m.stop_mel_token
        unfinished = torch.ones(ids.shape[0], dtype=torch.long)
        pkv = None
        codes, keep, lats, pos_trace = [], [], [], []
        orig = im.text_pos_embedding.get_fixed_embedding
        im.text_pos_embedding.get_fixed_embedding = lambda ind, dev: (pos_trace.append(int(ind)), orig(ind, dev))[1]
        try:
            for s in range(max_new):
                mask = torch.ones_like(ids)
                inp = im.prepare_inputs_for_generation(ids, past_key_values=pkv, attention_mask=mask, use_cache=True)
                out = im(**inp, return_dict=True, output_hidden_states=True)
                pkv = out.past_key_values
                logits = out.logits[:, -1, :].float()
                if len(keep) < want_logits:
                    keep.append(logits.clone())
                lats.append(m.final_norm(out.hidden_states[-1][:, -1]).float())
                scores = warp(ids, procs(ids, logits))
                if do_sample:
                    nxt = torch.multinomial(torch.softmax(scores, -1), 1).squeeze(1)
                else:
                    nxt = scores.argmax(-1)
                nxt = nxt * unfinished + pad * (1 - unfinished)
                codes.append(nxt)
                ids = torch.cat([ids, nxt[:, None]], -1)
                if not ignore_eos:
                    unfinished = unfinished * (nxt != eos).long()
                if unfinished.max() == 0 or ids.shape[1] >= P1 + max_new:
                    break
        finally:
            im.text_pos_embedding.get_fixed_embedding = orig
        return dict(codes=torch.stack(codes, 1), logits=torch.stack(keep) if keep else None,
                    latents=torch.stack(lats, 1), pos_trace=[0] + pos_trace, prefix_len=P1 - 1)


def load_reference_clvp():
    """The reference's CLVP class (tortoise_tts/models/clvp.py).  Its module imports the non-xformers Transformer, which needs
    `rotary_embedding_torch` (absent here, never called with use_xformers=True): stubbed."""
    load_reference()
    if "rotary_embedding_torch" not in sys.modules:
        m = types.ModuleType("rotary_embedding_torch")
        m.__spec__ = importlib.machinery.ModuleSpec("rotary_embedding_torch", None)
        m.RotaryEmbedding = type("RotaryEmbedding", (), {})
        m.broadcat = None
        sys.modules["rotary_embedding_torch"] = m
    return importlib.import_module("tortoise_tts.models.clvp")


def build_reference_clvp(sd, cfg):
    """Reference CLVP at cfg's size with the synthetic weights loaded strictly, in eval mode."""
    clvp = load_reference_clvp()
    m = clvp.CLVP(dim_text=cfg.dim, dim_speech=cfg.dim, dim_latent=cfg.dim, num_text_tokens=cfg.num_text_tokens,
                  text_enc_depth=cfg.depth, text_heads=cfg.heads, num_speech_tokens=cfg.num_speech_tokens,
                  speech_enc_depth=cfg.depth, speech_heads=cfg.heads, use_xformers=True)
    m.load_state_dict(sd, strict=True)
    m.eval()
    return m
