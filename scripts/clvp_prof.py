"""One CLVP scoring call of the C2 shape (16 x 500 codes, 64 text tokens, depth 20) -- for ncu launch lists."""
import os
import sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tortoise_tts_b200.clvp import CLVPB200  # noqa: E402
from tortoise_tts_b200.synthetic import ClvpConfig, make_clvp_inputs, make_clvp_state_dict  # noqa: E402
cfg = ClvpConfig()
sd = make_clvp_state_dict(cfg, seed=5)
text, codes = make_clvp_inputs(cfg, 16, 64, 500, seed=6)
m = CLVPB200(sd, depth=cfg.depth, max_seqs=16, max_text_len=64, max_speech_len=500)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    s = m(text.cuda(), codes.cuda())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    s = m(text.cuda(), codes.cuda())
e1.record()
torch.cuda.synchronize()
print("clvp ms/call", e0.elapsed_time(e1) / 5, s[:4].tolist())
