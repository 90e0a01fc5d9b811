"""B200-native IQ front end + RxVFO channelizer bank (SDR++Brown hot path).

The product is the C-ABI library ``libb200dsp.so`` (``include/b200dsp.h``) built from
``sdrplusplusbrown_b200/csrc``; this package is the thin Python host binding used by
tests and bench.py.  There is no CPU fallback: importing :mod:`.binding` fails loudly
when the CUDA library is missing.
"""
__all__ = ["binding", "synth", "configs"]
