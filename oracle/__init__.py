"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's UnifiedVoice autoregressive sampler
(e-c-k-e-r/tortoise-tts, `tortoise_tts/models/unified_voice.py` +
`tortoise_tts/models/stream_generator.py` + the HF GPT-2 block arithmetic the
reference calls into).  Nothing in the product package (`tortoise_tts_b200/`)
may import this package: only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` do, and only as the
checker / the timed CPU baseline.

Parity pin: the reference ships no tests or golden vectors for this path
(SURVEY.md section 4), so the restatement is pinned against the reference's own
model classes imported from /root/reference (oracle/ref_shim.py, runs only in
the build container) -- see oracle/make_golden.py and tests/golden/.
"""
