"""Host-side mirror of the reference's AR model interface, on top of the C ABI (include/tts_ar.h).

`UnifiedVoiceB200` keeps the method names, argument meaning and error behaviour of the reference's
`UnifiedVoice` for the autoregressive path (reference tortoise_tts/models/unified_voice.py):
    inference_speech      :632-668      compute_embeddings :614-630
    get_generator         :670-679      forward(..., return_latent=True) :544-600
so `tortoise_tts/inference.py:266-285, 334-346, 371-379` and `train.py:119-129` can call it unchanged
(see INTEGRATION.md for `patch_reference_module`).  PyTorch is used only for device memory and the
current CUDA stream; all arithmetic happens in libtts_ar.so.
"""
import ctypes as C

import torch

from . import _lib


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class ARGeometry:
    """Defaults = UnifiedVoice.__init__ defaults (unified_voice.py:335-354)."""

    def __init__(self, layers=30, model_dim=1024, heads=16, max_text_tokens=402, max_mel_tokens=604,
                 max_conditioning_inputs=2, number_text_tokens=255, start_text_token=255, stop_text_token=0,
                 number_mel_codes=8194, start_mel_token=8192, stop_mel_token=8193):
        self.layers, self.model_dim, self.heads = layers, model_dim, heads
        self.max_text_tokens, self.max_mel_tokens = max_text_tokens, max_mel_tokens
        self.max_conditioning_inputs = max_conditioning_inputs
        self.number_text_tokens, self.start_text_token, self.stop_text_token = number_text_tokens, start_text_token, stop_text_token
        self.number_mel_codes, self.start_mel_token, self.stop_mel_token = number_mel_codes, start_mel_token, stop_mel_token

    @property
    def n_mel_pos(self):
        return self.max_mel_tokens + 2 + self.max_conditioning_inputs

    @property
    def n_text_pos(self):
        return self.max_text_tokens + 2


class UnifiedVoiceB200:
    def __init__(self, state_dict, geometry=None, device="cuda:0", max_utts=1, max_seqs=16, max_text_len=None,
                 max_new_tokens=None):
        g = geometry or ARGeometry()
        self.geometry = g
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("UnifiedVoiceB200 runs on a CUDA (sm_100a) device only; there is no CPU path")
        self.lib = _lib.load()
        self.max_mel_tokens = g.max_mel_tokens
        self.start_mel_token, self.stop_mel_token = g.start_mel_token, g.stop_mel_token
        self.start_text_token, self.stop_text_token = g.start_text_token, g.stop_text_token
        cfg = _lib.TtsArConfig()
        cfg.struct_size = C.sizeof(cfg)
        cfg.device = self.device.index or 0
        cfg.n_layers, cfg.model_dim, cfg.n_heads = g.layers, g.model_dim, g.heads
        cfg.n_mel_codes, cfg.n_text_rows = g.number_mel_codes, g.number_text_tokens + 1
        cfg.n_mel_pos, cfg.n_text_pos = g.n_mel_pos, g.n_text_pos
        cfg.start_text_token, cfg.stop_text_token = g.start_text_token, g.stop_text_token
        cfg.start_mel_token, cfg.stop_mel_token = g.start_mel_token, g.stop_mel_token
        cfg.max_utts, cfg.max_seqs = max_utts, max_seqs
        cfg.max_text_len = g.max_text_tokens if max_text_len is None else max_text_len
        cfg.max_new_tokens = (g.max_mel_tokens - 1) if max_new_tokens is None else max_new_tokens
        self.cfg = cfg
        self.handle = C.c_void_p(0)
        _lib.check(self.lib, None, self.lib.tts_ar_create(C.byref(cfg), C.byref(self.handle)), "tts_ar_create")
        self._embed_state = None
        self.load_state_dict(state_dict)

    # ---- lifetime -------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "handle", None) and self.handle.value:
            try:
                self._close_stream()
            except Exception:
                pass
            self.lib.tts_ar_destroy(self.handle)
            self.handle = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def module_weights(module):
        """{reference state_dict key: tensor} for the AR path, collected by walking `named_modules()` and READING the
        `.weight` / `.bias` attributes.  With the reference's LoRA active (`torch.nn.utils.parametrize`, reference
        lora.py:163-195) `state_dict()` holds `...parametrizations.weight.original` (the un-merged base weight) while
        the attribute returns the merged value W + (B @ A) * scale; a swapped-in `LoRAConv1d` / `LoRALinear`
        (lora.py:21-140) keeps `.weight` as the base and carries `lora_A`, `lora_B`, `scaling`, `merged`: those are merged
        here the way the module's own `forward` applies them."""
        import torch.nn as nn
        out = {}
        for name, mod in module.named_modules():
            if name.startswith("inference_model") or ".parametrizations" in name:
                continue   # inference_model shares gpt / embeddings / heads (unified_voice.py:453-487): same tensors
            for attr in ("weight", "bias"):
                t = getattr(mod, attr, None)
                if not isinstance(t, torch.Tensor):
                    continue
                t = t.detach()
                if attr == "weight" and hasattr(mod, "lora_A") and hasattr(mod, "lora_B") and not getattr(mod, "merged", False) \
                        and getattr(mod, "enabled", True):
                    a, b = mod.lora_A.detach(), mod.lora_B.detach()
                    scale = float(getattr(mod, "scaling", getattr(mod, "scale", 1.0)))
                    delta = (b @ a) * scale if b.shape[0] == t.shape[0] else ((b @ a) * scale).t()
                    if delta.shape == t.shape:
                        t = t + delta.to(t.dtype)
                out[f"{name}.{attr}" if name else attr] = t
        return out

    @classmethod
    def from_module(cls, module, **kw):
        """Snapshot a reference `UnifiedVoice` nn.Module; LoRA-merged values are taken (see `module_weights`)."""
        g = ARGeometry(layers=module.layers, model_dim=module.model_dim, heads=module.heads,
                       max_text_tokens=module.max_text_tokens, max_mel_tokens=module.max_mel_tokens,
                       max_conditioning_inputs=module.max_conditioning_inputs,
                       number_text_tokens=module.number_text_tokens, start_text_token=module.start_text_token,
                       stop_text_token=module.stop_text_token, number_mel_codes=module.number_mel_codes,
                       start_mel_token=module.start_mel_token, stop_mel_token=module.stop_mel_token)
        return cls(cls.module_weights(module), geometry=g, **kw)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _check(self, rc, what):
        _lib.check(self.lib, self.handle, rc, what)

    def set_option(self, name, value):
        """Explicit per-handle option (include/tts_ar.h tts_ar_set_option)."""
        self._check(self.lib.tts_ar_set_option(self.handle, name.encode(), int(value)), f"tts_ar_set_option({name})")

    def load_state_dict(self, sd):
        """Re-snapshot the weights (call after the module's load_state_dict or a LoRA toggle)."""
        g = self.geometry
        dt = torch.bfloat16 if all(v.dtype == torch.bfloat16 for k, v in sd.items() if k.startswith("gpt.h.0.")) else torch.float32
        keep = []

        def dev(name):
            t = sd[name].detach().to(device=self.device, dtype=dt).contiguous()
            keep.append(t)
            return C.c_void_p(t.data_ptr())

        layers = (_lib.TtsArLayerWeights * g.layers)()
        for i in range(g.layers):
            p = f"gpt.h.{i}."
            lw = layers[i]
            lw.ln_1_w, lw.ln_1_b = dev(p + "ln_1.weight"), dev(p + "ln_1.bias")
            lw.c_attn_w, lw.c_attn_b = dev(p + "attn.c_attn.weight"), dev(p + "attn.c_attn.bias")
            lw.attn_proj_w, lw.attn_proj_b = dev(p + "attn.c_proj.weight"), dev(p + "attn.c_proj.bias")
            lw.ln_2_w, lw.ln_2_b = dev(p + "ln_2.weight"), dev(p + "ln_2.bias")
            lw.c_fc_w, lw.c_fc_b = dev(p + "mlp.c_fc.weight"), dev(p + "mlp.c_fc.bias")
            lw.mlp_proj_w, lw.mlp_proj_b = dev(p + "mlp.c_proj.weight"), dev(p + "mlp.c_proj.bias")
        w = _lib.TtsArWeights()
        w.struct_size = C.sizeof(w)
        w.dtype = 1 if dt == torch.bfloat16 else 0
        w.n_layers = g.layers
        w.layers = layers
        w.ln_f_w, w.ln_f_b = dev("gpt.ln_f.weight"), dev("gpt.ln_f.bias")
        w.final_norm_w, w.final_norm_b = dev("final_norm.weight"), dev("final_norm.bias")
        w.mel_head_w, w.mel_head_b = dev("mel_head.weight"), dev("mel_head.bias")
        w.mel_embedding = dev("mel_embedding.weight")
        w.mel_pos_embedding = dev("mel_pos_embedding.emb.weight")
        w.text_embedding = dev("text_embedding.weight")
        w.text_pos_embedding = dev("text_pos_embedding.emb.weight")
        torch.cuda.synchronize(self.device)
        self._check(self.lib.tts_ar_load_weights(self.handle, C.byref(w), self._stream()), "tts_ar_load_weights")
        # the conditioning-latent encoder, when the state_dict carries it (the AR path itself does not need it)
        self.has_conditioning_encoder = "conditioning_encoder.init.weight" in sd
        if self.has_conditioning_encoder:
            n_blocks = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("conditioning_encoder.attn."))
            blocks = (_lib.TtsArCondBlockWeights * n_blocks)()
            for i in range(n_blocks):
                a = f"conditioning_encoder.attn.{i}."
                b = blocks[i]
                b.norm_w, b.norm_b = dev(a + "norm.weight"), dev(a + "norm.bias")
                b.qkv_w, b.qkv_b = dev(a + "qkv.weight"), dev(a + "qkv.bias")
                b.proj_w, b.proj_b = dev(a + "proj_out.weight"), dev(a + "proj_out.bias")
            cw = _lib.TtsArCondWeights()
            cw.struct_size, cw.dtype, cw.n_blocks = C.sizeof(cw), w.dtype, n_blocks
            cw.spec_dim = sd["conditioning_encoder.init.weight"].shape[1]
            cw.init_w, cw.init_b = dev("conditioning_encoder.init.weight"), dev("conditioning_encoder.init.bias")
            cw.blocks = blocks
            torch.cuda.synchronize(self.device)
            self._check(self.lib.tts_ar_load_conditioning_weights(self.handle, C.byref(cw), self._stream()),
                        "tts_ar_load_conditioning_weights")
        torch.cuda.synchronize(self.device)
        del keep

    def mel_spectrogram(self, inp, mel_norms=None):
        """TorchMelSpectrogram.forward (tortoise_tts/models/arch_utils.py:361-395): [b, n] or [b, 1, n] audio at 22.05 kHz ->
        [b, 80, 1 + n // 256] normalised log-mel (what emb/mel.py:50-65 feeds to get_conditioning).  `mel_norms`: the 80 values
        of the reference's data/models/mel_norms.pth (TorchMelSpectrogram.mel_norms), or None."""
        x = inp.squeeze(1) if inp.dim() == 3 else inp
        if x.dim() != 2:
            raise ValueError("audio must be [b, n] or [b, 1, n]")
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        norms = None if mel_norms is None else mel_norms.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty(x.shape[0], 80, 1 + x.shape[1] // 256, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.tts_ar_mel_spectrogram(self.handle, _ptr(x), x.shape[0], x.shape[1], _ptr(norms), _ptr(out), self._stream()),
                        "tts_ar_mel_spectrogram")
        return out

    def get_conditioning(self, speech_conditioning_input):
        """UnifiedVoice.get_conditioning (unified_voice.py:535-542): [b, n_clips, 80, frames] (or [b, 80, frames]) mel
        spectrograms -> [b, 1024] conditioning latents (mean over the clips of column 0 of the conditioning encoder)."""
        if not getattr(self, "has_conditioning_encoder", False):
            raise RuntimeError("the state_dict this model was built from has no conditioning_encoder.* entries")
        x = speech_conditioning_input
        x = x.unsqueeze(1) if x.dim() == 3 else x
        if x.dim() != 4:
            raise ValueError("speech_conditioning_input must be [b, n_clips, spec, frames] or [b, spec, frames]")
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty(x.shape[0], self.geometry.model_dim, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            for b in range(x.shape[0]):
                self._check(self.lib.tts_ar_get_conditioning(self.handle, _ptr(x[b]), x.shape[1], x.shape[3], _ptr(out[b]), self._stream()),
                            "tts_ar_get_conditioning")
            self.sync_stats()
        return out

    # ---- helpers ----------------------------------------------------------------------------------------
    def _sampling(self, kw):
        kw = dict(kw)
        num_beams = kw.pop("num_beams", 1)
        if num_beams is not None and num_beams > 1:
            # the reference's beam modes call HF methods that no longer exist
            raise NotImplementedError("beam search is not supported (nor by the reference on current transformers)")
        s = _lib.TtsArSampling()
        s.struct_size = C.sizeof(s)
        s.do_sample = 1 if kw.pop("do_sample", False) else 0
        t = kw.pop("temperature", 1.0)
        s.temperature = 1.0 if t is None else float(t)
        k = kw.pop("top_k", 50)                      # HF GenerationConfig default
        s.top_k = 0 if k is None else int(k)
        p = kw.pop("top_p", 1.0)
        s.top_p = 1.0 if p is None else float(p)
        r = kw.pop("repetition_penalty", 1.0)
        s.repetition_penalty = 1.0 if r is None else float(r)
        s.seed = int(kw.pop("seed", 0))              # reference generate(seed=0), stream_generator.py:223
        s.seq_index_base = int(kw.pop("seq_index_base", 0))
        idx = kw.pop("seq_indices", None)          # global index per local sequence (non-contiguous shards, shard.py)
        if idx is not None:
            self._seq_indices = torch.as_tensor(idx, dtype=torch.int64).to(self.device).contiguous()   # kept alive for the call
            s.seq_indices = self._seq_indices.data_ptr()
        s.ignore_eos = 1 if kw.pop("ignore_eos", False) else 0
        # reference: LogitsProcessorList([TypicalLogitsWarper(mass=typical_mass)]) if typical_sampling (:659)
        tm = kw.pop("typical_mass", None)
        s.typical_mass = 0.0 if tm is None else float(tm)
        for ignored in ("length_penalty", "output_attentions", "output_hidden_states", "return_dict_in_generate",
                        "logits_processor", "bos_token_id", "pad_token_id", "eos_token_id", "use_cache"):
            kw.pop(ignored, None)
        if kw:
            raise ValueError(f"unsupported generate kwargs: {sorted(kw)}")
        if s.do_sample and s.temperature <= 0:
            raise ValueError("`temperature` has to be a strictly positive float")
        return s

    def _prefill(self, latent, text_inputs, n_cand, text_lengths=None):
        if latent.dim() != 2 or latent.shape[1] != self.geometry.model_dim:
            raise ValueError("speech_conditioning_latent must be [b, 1024]")
        if text_inputs.dim() != 2 or text_inputs.shape[0] != latent.shape[0]:
            raise ValueError("text_inputs must be [b, t] with the latent's batch size")
        U, T = text_inputs.shape
        lat = latent.detach().to(device=self.device, dtype=torch.float32).contiguous()
        txt = text_inputs.detach().to(device=self.device, dtype=torch.int32).contiguous()
        # Reference semantics: a padded batch is NOT masked, pads are real stop_text tokens (quirk Q8), so the
        # default is the full row; `text_lengths` is the ragged extension.
        lens = [T] * U if text_lengths is None else [int(x) for x in text_lengths]
        arr = (C.c_int32 * U)(*lens)
        self._check(self.lib.tts_ar_prefill(self.handle, _ptr(lat), _ptr(txt), max(T, 1), arr, U, int(n_cand), self._stream()),
                    "tts_ar_prefill")
        return U, lens

    def sync_stats(self):
        st = _lib.TtsArStats()
        self._check(self.lib.tts_ar_sync_result(self.handle, C.byref(st), self._stream()), "tts_ar_sync_result")
        return st

    def fix_autoregressive_codes(self, codes, calm_token=83, strict=True):
        """The epilogue TTS.inference / train.run_eval run on the sampled codes, on the device and in place
        (reference inference.py:353-366, 381-389; train.py:135-147, 161-169): in every row the stop token and everything
        after it become `calm_token`, the last three codes become 45, 45, 248.  Returns (codes, trim) where trim is the
        index at which the reference cuts the latents (first index of row 0 with more than 8 consecutive calm tokens
        behind it; the row length if there is none).  A row without a stop token makes the reference raise (`.min()`
        of an empty tensor, inference.py:355): with strict=True so does this, with strict=False the row is left as is."""
        if codes.dim() != 2 or codes.dtype != torch.int64 or not codes.is_cuda:
            raise ValueError("codes must be a CUDA int64 tensor [rows, len]")
        if codes.stride(1) != 1:
            raise ValueError("codes rows must be contiguous")
        rows, n = codes.shape
        first = torch.empty(rows, device=codes.device, dtype=torch.int32)
        trim = torch.empty(rows, device=codes.device, dtype=torch.int32)
        self._check(self.lib.tts_ar_fixup_codes(self.handle, _ptr(codes), rows, n, codes.stride(0), int(calm_token), _ptr(first),
                                                _ptr(trim), self._stream()), "tts_ar_fixup_codes")
        first, trim = first.cpu(), trim.cpu()          # one device -> host read for both results (the reference: one per row)
        if strict and rows and bool((first < 0).any()):
            raise RuntimeError("min(): Expected reduction dim to be specified for input.numel() == 0. "
                               "(a row without a stop token: reference inference.py:355)")
        return codes, (int(trim[0]) if rows else n)

    def logits_debug(self):
        rows = C.c_int32(0)
        out = torch.empty(self.cfg.max_seqs, self.geometry.number_mel_codes, device=self.device, dtype=torch.float32)
        self._check(self.lib.tts_ar_logits_debug(self.handle, _ptr(out), C.byref(rows), self._stream()), "tts_ar_logits_debug")
        return out[: rows.value]

    def debug_buffer(self, name, shape, dtype=torch.float32):
        out = torch.empty(shape, device=self.device, dtype=dtype)
        self._check(self.lib.tts_ar_debug_copy(self.handle, name.encode(), _ptr(out), out.numel() * out.element_size(),
                                               self._stream()), "tts_ar_debug_copy")
        torch.cuda.synchronize(self.device)
        return out

    # ---- reference surface ----------------------------------------------------------------------------------
    def inference_speech(self, speech_conditioning_latent, text_inputs, input_tokens=None, num_return_sequences=1,
                         max_generate_length=None, typical_sampling=False, typical_mass=.9, kv_cache=True,
                         text_lengths=None, return_latents=False, forced_tokens=None, return_choice=False,
                         **hf_generate_kwargs):
        if input_tokens is not None:
            raise NotImplementedError("input_tokens (continuation) is not supported")
        if typical_sampling:
            if not 0.0 < float(typical_mass) < 1.0:
                raise ValueError("typical_mass must be in (0, 1)")
            hf_generate_kwargs = dict(hf_generate_kwargs, typical_mass=typical_mass)
        s = self._sampling(hf_generate_kwargs)
        max_new = (self.max_mel_tokens - 1) if max_generate_length is None else int(max_generate_length)
        U, _ = self._prefill(speech_conditioning_latent, text_inputs, num_return_sequences, text_lengths)
        B = U * num_return_sequences
        codes = torch.full((B, max_new), self.stop_mel_token, dtype=torch.int64, device=self.device)
        lat = torch.zeros((B, max_new, self.geometry.model_dim), dtype=torch.float32, device=self.device) if return_latents else None
        forced = forced_tokens.to(device=self.device, dtype=torch.int32).contiguous() if forced_tokens is not None else None
        choice = torch.zeros((B, max_new), dtype=torch.int32, device=self.device) if return_choice else None
        self._check(self.lib.tts_ar_generate(self.handle, C.byref(s), max_new, _ptr(codes), _ptr(lat), _ptr(forced),
                                             _ptr(choice), self._stream()), "tts_ar_generate")
        st = self.sync_stats()
        self.last_stats = st
        out = codes[:, : st.steps_run].to(text_inputs.device) if text_inputs.is_cuda else codes[:, : st.steps_run].cpu()
        extras = []
        if return_latents:
            extras.append(lat[:, : st.steps_run])
        if return_choice:
            extras.append(choice[:, : st.steps_run])
        return (out, *extras) if extras else out

    def compute_embeddings(self, cond_latents, text_inputs, kv_cache=True):
        U, T = text_inputs.shape
        self._embed_state = (cond_latents, text_inputs)
        embs = torch.full((U, T + 3 + 1), 1, dtype=torch.long, device=text_inputs.device)
        embs[:, -1] = self.start_mel_token
        return embs

    def get_generator(self, inputs, max_length=500, chunk=None, **hf_generate_kwargs):
        """Yields (next_tokens [B] int64, latent [B, 1024]) per step like sample_stream (stream_generator.py:1172).
        `max_length` counts the prompt (quirk Q9).  ONE persistent launch on a side stream runs the whole sample loop and
        publishes every step through a flag ring in mapped pinned memory (include/tts_ar.h tts_ar_stream_*): per yielded
        step the host only polls a flag and slices the device buffers -- no launch, allocation or stream synchronisation.
        (`chunk` is accepted for backward compatibility and ignored.)"""
        if self._embed_state is None:
            raise AssertionError("compute_embeddings must be called first")   # ref: assert cached_mel_emb is not None
        n_cand = int(hf_generate_kwargs.pop("num_return_sequences", 1))
        s = self._sampling(hf_generate_kwargs)
        self._close_stream()
        latent, text = self._embed_state
        U, _ = self._prefill(latent, text, n_cand)
        B = U * n_cand
        total = int(max_length) - inputs.shape[1]
        total = min(total, self.cfg.max_new_tokens)
        if total < 1:
            return
        codes = torch.full((B, total), self.stop_mel_token, dtype=torch.int64, device=self.device)
        lat = torch.empty((B, total, self.geometry.model_dim), dtype=torch.float32, device=self.device)
        if getattr(self, "_side_stream", None) is None:
            self._side_stream = torch.cuda.Stream(device=self.device)
        side = self._side_stream
        side.wait_stream(torch.cuda.current_stream(self.device))          # prompt pass and buffer fills precede the loop
        sp = C.c_void_p(side.cuda_stream)
        self._check(self.lib.tts_ar_stream_begin(self.handle, C.byref(s), total, _ptr(codes), _ptr(lat), sp), "tts_ar_stream_begin")
        self._open_stream = (sp, codes, lat)
        try:
            for k in range(total):
                r = self.lib.tts_ar_stream_poll(self.handle, k, -1)
                if r < 0:
                    self._check(r, "tts_ar_stream_poll")
                if r == 2:
                    break
                yield codes[:, k], lat[:, k]
        finally:
            self._close_stream()

    def _close_stream(self):
        st = getattr(self, "_open_stream", None)
        if st is None:
            return
        self._open_stream = None
        stats = _lib.TtsArStats()
        self._check(self.lib.tts_ar_stream_end(self.handle, C.byref(stats), st[0]), "tts_ar_stream_end")
        self.last_stats = stats

    def forward(self, speech_conditioning_latent, text_inputs, text_lengths, mel_codes, wav_lengths, types=None,
                text_first=True, raw_mels=None, return_attentions=False, return_latent=False, clip_inputs=True, out=None,
                ragged_text=False):
        if not return_latent or not text_first or raw_mels is not None or types is not None or return_attentions:
            raise NotImplementedError("only forward(..., return_latent=True, text_first=True) is on the AR inference path")
        n = mel_codes.shape[0]
        # the reference caller passes lengths as [n] or [n, 1] (inference.py:371-379: wav_lengths.repeat(candidates, 1));
        # set_mel_padding / the clipping below are per row (unified_voice.py:494-506, 565-571)
        wav_lengths = torch.as_tensor(wav_lengths).reshape(-1)
        text_lengths = torch.as_tensor(text_lengths).reshape(-1)
        if wav_lengths.numel() == 1:
            wav_lengths = wav_lengths.expand(n)
        if text_lengths.numel() == 1:
            text_lengths = text_lengths.expand(n)
        if wav_lengths.numel() != n or text_lengths.numel() != n:
            raise ValueError("text_lengths / wav_lengths must have one entry per row of mel_codes")
        if clip_inputs:
            max_text_len = int(text_lengths.max())
            text_inputs = text_inputs[:, :max_text_len]
            max_mel_len = int(wav_lengths.max()) // 1024
            mel_codes = mel_codes[:, :max_mel_len]
        n, M = mel_codes.shape
        T = text_inputs.shape[1]
        # set_mel_padding (unified_voice.py:494-506): positions past wav_len//1024 + 1 become stop tokens
        codes = mel_codes.detach().clone().to(self.device)
        mel_lengths = torch.div(wav_lengths.to(self.device), 1024, rounding_mode="trunc")
        idx = torch.arange(M, device=self.device)[None, :]
        codes = torch.where(idx >= (mel_lengths[:, None] + 1), torch.full_like(codes, self.stop_mel_token), codes)
        assert codes.shape == (n, M)
        codes = codes.to(torch.int32).contiguous()
        lat = speech_conditioning_latent.detach().to(device=self.device, dtype=torch.float32).contiguous()
        txt = text_inputs.detach().to(device=self.device, dtype=torch.int32).contiguous()
        if out is None:
            out = torch.zeros((n, M, self.geometry.model_dim), dtype=torch.float32, device=self.device)
        elif out.shape != (n, M, self.geometry.model_dim) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != self.device:
            raise ValueError("out must be a contiguous float32 [n, M, 1024] tensor on the model's device")   # e.g. this rank's slice of an all-gather buffer
        # reference semantics: padded text is NOT masked (quirk Q8) -> every row runs at the full width T; ragged_text=True
        # is the extension that runs each row at its own text_lengths[i] (what a ragged inference_speech batch produced)
        tl = (C.c_int32 * n)(*([int(x) for x in text_lengths] if ragged_text else [T] * n))
        cl = (C.c_int32 * n)(*([M] * n))
        self._check(self.lib.tts_ar_forward_latents(self.handle, _ptr(lat), _ptr(txt), max(T, 1), tl, _ptr(codes), max(M, 1),
                                                    cl, n, _ptr(out), self._stream()), "tts_ar_forward_latents")
        self.sync_stats()
        return out

    __call__ = forward

    def inference_speech_host(self, latent_cpu, text_cpu, num_return_sequences=1, max_generate_length=None, **kw):
        """End-to-end entry on HOST tensors (pinned staging, H2D, prefill, sample loop, D2H inside the C call)."""
        s = self._sampling(kw)
        max_new = (self.max_mel_tokens - 1) if max_generate_length is None else int(max_generate_length)
        lat = latent_cpu.detach().to(dtype=torch.float32, device="cpu").contiguous()
        txt = text_cpu.detach().to(dtype=torch.int32, device="cpu").contiguous()
        U, T = txt.shape
        out = torch.full((U * num_return_sequences, max_new), self.stop_mel_token, dtype=torch.int64)
        lens = (C.c_int32 * U)(*([T] * U))
        steps = C.c_int32(0)
        self._check(self.lib.tts_ar_inference_speech_host(self.handle, _ptr(lat), _ptr(txt), T, lens, U,
                                                          int(num_return_sequences), C.byref(s), max_new, _ptr(out),
                                                          C.byref(steps), self._stream()), "tts_ar_inference_speech_host")
        return out[:, : steps.value]


def patch_reference_module(module, **kw):
    """Route a reference `UnifiedVoice` instance's AR entry points through the B200 path, in place.
    (`engine.module` in tortoise_tts/inference.py:185-202 is such an instance.)"""
    impl = UnifiedVoiceB200.from_module(module, device=next(module.parameters()).device, **kw)
    module._b200 = impl
    module.inference_speech = impl.inference_speech
    module.compute_embeddings = impl.compute_embeddings
    module.get_generator = impl.get_generator
    if impl.has_conditioning_encoder:
        module.get_conditioning = impl.get_conditioning
    orig_forward = module.forward

    def forward(*a, **k):
        if k.get("return_latent", False):
            return impl.forward(*a, **k)
        return orig_forward(*a, **k)

    module.forward = forward
    return module
