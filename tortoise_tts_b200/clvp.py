"""CLVPB200 -- the reference's CLVP candidate scorer (tortoise_tts/models/clvp.py) on the B200 library, behind the
reference's own call: `scores = clvp(text_tokens.repeat(n, 1), codes, return_loss=False)` (tortoise_tts/inference.py:392-396;
the caller keeps `torch.topk(scores, k=candidates)`).

Only the inference call is mirrored (eval mode, use_xformers=True, return_loss=False): the contrastive loss
(`return_loss=True`) and the training-time token masking belong to training and raise NotImplementedError here.
No CPU / PyTorch fallback: the constructor raises when libtts_ar.so is missing or the device is not an sm_100 GPU.
"""
import ctypes as C

import torch

from . import _lib


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class CLVPB200:
    def __init__(self, state_dict, depth=20, dim=768, heads=12, num_text_tokens=256, num_speech_tokens=8192, device=None,
                 max_seqs=16, max_text_len=402, max_speech_len=608):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError("CLVPB200 needs a CUDA device (sm_100); there is no CPU fallback")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.depth, self.dim, self.heads = depth, dim, heads
        self.num_text_tokens, self.num_speech_tokens = num_text_tokens, num_speech_tokens
        self.max_seqs, self.max_text_len, self.max_speech_len = max_seqs, max_text_len, max_speech_len
        cfg = _lib.TtsClvpConfig(C.sizeof(_lib.TtsClvpConfig), self.device.index, dim, heads, depth, num_text_tokens,
                                 num_speech_tokens, max_seqs, max_text_len, max_speech_len)
        self.handle = C.c_void_p(0)
        rc = self.lib.tts_ar_clvp_create(C.byref(cfg), C.byref(self.handle))
        if rc != 0:
            self._raise(rc, "tts_ar_clvp_create", None)
        self.training = False
        self.load_state_dict(state_dict)

    # ---- construction from the reference module -------------------------------------------------------------------------------
    @classmethod
    def from_module(cls, module, **kw):
        """Snapshot a reference `CLVP` nn.Module (built with use_xformers=True, the reference default, clvp.py:45)."""
        if not getattr(module, "xformers", True):
            raise ValueError("CLVPB200 mirrors the use_xformers=True encoders only")
        sd = {k: v.detach() for k, v in module.state_dict().items()}
        depth = 1 + max(int(k.split(".")[4]) for k in sd if k.startswith("text_transformer.transformer.attn_layers.layers.")) // 2
        dim = sd["text_emb.weight"].shape[1]
        heads = sd["text_transformer.transformer.attn_layers.layers.0.1.wrap.to_q.weight"].shape[0] // 64
        return cls(sd, depth=depth, dim=dim, heads=heads, num_text_tokens=sd["text_emb.weight"].shape[0],
                   num_speech_tokens=sd["speech_emb.weight"].shape[0], **kw)

    def _raise(self, rc, what, handle):
        msg = self.lib.tts_ar_clvp_last_error(handle)
        text = f"{what}: {_lib.STATUS.get(rc, rc)}: {msg.decode() if msg else ''}"
        raise (ValueError if rc in (-1, -2) else RuntimeError)(text)

    def _check(self, rc, what):
        if rc != 0:
            self._raise(rc, what, self.handle)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def load_state_dict(self, sd):
        """Re-snapshot the weights (reference state_dict keys of CLVP, see include/tts_ar.h TtsClvpLayerWeights)."""
        dt = torch.bfloat16 if sd["text_emb.weight"].dtype == torch.bfloat16 else torch.float32
        keep = []

        def dev(key):
            t = sd[key].detach().to(device=self.device, dtype=dt).contiguous()
            keep.append(t)
            return t.data_ptr()

        w = _lib.TtsClvpWeights()
        w.struct_size, w.dtype, w.depth = C.sizeof(_lib.TtsClvpWeights), 1 if dt == torch.bfloat16 else 0, self.depth
        arrays = []
        for side, ew in (("text", w.text), ("speech", w.speech)):
            p = f"{side}_transformer.transformer."
            layers = (_lib.TtsClvpLayerWeights * self.depth)()
            for i in range(self.depth):
                a, f = f"{p}attn_layers.layers.{2 * i}.", f"{p}attn_layers.layers.{2 * i + 1}."
                L = layers[i]
                L.attn_norm_g = dev(a + "0.0.g")
                L.to_q_w, L.to_k_w, L.to_v_w = (dev(a + f"1.wrap.{n}.weight") for n in ("to_q", "to_k", "to_v"))
                L.to_out_w, L.to_out_b = dev(a + "1.wrap.to_out.weight"), dev(a + "1.wrap.to_out.bias")
                L.ff_norm_g = dev(f + "0.0.g")
                L.ff_proj_w, L.ff_proj_b = dev(f + "1.wrap.net.0.proj.weight"), dev(f + "1.wrap.net.0.proj.bias")
                L.ff_out_w, L.ff_out_b = dev(f + "1.wrap.net.3.weight"), dev(f + "1.wrap.net.3.bias")
            arrays.append(layers)
            ew.emb = dev(f"{side}_emb.weight")
            ew.layers = layers
            ew.norm_w, ew.norm_b = dev(p + "norm.weight"), dev(p + "norm.bias")
            ew.to_latent_w = dev(f"to_{side}_latent.weight")
        w.temperature = dev("temperature")
        with torch.cuda.device(self.device):
            self._check(self.lib.tts_ar_clvp_load_weights(self.handle, C.byref(w), self._stream()), "tts_ar_clvp_load_weights")
            torch.cuda.current_stream(self.device).synchronize()      # the staged copies in `keep` may go now
        del keep, arrays

    # ---- the reference's call --------------------------------------------------------------------------------------------------
    def forward(self, text, speech_tokens, return_loss=False, return_latents=False):
        """CLVP.forward(text, speech_tokens, return_loss=False) (clvp.py:100-133): text [n, T] (or [1, T]) and speech_tokens
        [n, L] integer tensors -> sim [n] float32 on the device.  `return_latents=True` (not in the reference) also returns the
        L2-normalised speech latents [n, dim] (parity hook)."""
        if return_loss:
            raise NotImplementedError("CLVPB200 mirrors the inference call (return_loss=False) only")
        if text.dim() != 2 or speech_tokens.dim() != 2:
            raise ValueError("text and speech_tokens must be [batch, length]")
        n = speech_tokens.shape[0]
        if text.shape[0] not in (1, n):
            raise ValueError("text must have one row, or one row per speech_tokens row")
        with torch.cuda.device(self.device):
            t32 = text.to(device=self.device, dtype=torch.int32).contiguous()
            c64 = speech_tokens.to(device=self.device, dtype=torch.int64)
            if c64.stride(-1) != 1:
                c64 = c64.contiguous()
            scores = torch.empty(n, dtype=torch.float32, device=self.device)
            lat = torch.empty(n, self.dim, dtype=torch.float32, device=self.device) if return_latents else None
            self._check(self.lib.tts_ar_clvp_score(self.handle, _ptr(t32), t32.shape[0], t32.shape[1], t32.stride(0), _ptr(c64), n,
                                                   c64.shape[1], c64.stride(0), _ptr(scores), _ptr(lat), self._stream()),
                        "tts_ar_clvp_score")
            launches = C.c_int64(0)
            self._check(self.lib.tts_ar_clvp_sync(self.handle, C.byref(launches), self._stream()), "tts_ar_clvp_sync")
            self.kernel_launches = launches.value
        return (scores, lat) if return_latents else scores

    __call__ = forward

    def topk(self, scores, k):
        """`torch.topk(scores, k=k).indices` (reference inference.py:395) on the device: int64 [k], best first."""
        scores = scores.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty(int(k), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.tts_ar_clvp_topk(self.handle, _ptr(scores), scores.numel(), int(k), _ptr(out), self._stream()),
                        "tts_ar_clvp_topk")
        return out

    def eval(self):
        return self

    def to(self, *a, **k):      # `clvp.to("cpu")` / auto_unload in the reference (inference.py:226, 393): weights stay resident
        return self

    def close(self):
        if getattr(self, "handle", None) and self.handle.value:
            self.lib.tts_ar_clvp_destroy(self.handle)
            self.handle = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
