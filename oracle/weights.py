"""Synthetic weights / inputs for the oracle and the tests (TEST INFRASTRUCTURE, see oracle/__init__.py).
The generator itself lives in the product package (tortoise_tts_b200/synthetic.py: pure data generation, shared
with bench.py); it is re-exported here so tests read `oracle.weights`."""
from tortoise_tts_b200.synthetic import ARConfig, make_inputs, make_state_dict  # noqa: F401
