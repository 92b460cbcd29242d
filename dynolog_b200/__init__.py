"""dynolog_b200: B200-native decode + aggregate of hbt perf_event ring-buffer images.

The product is the C-ABI shared library ``dynolog_b200/libhbtgpu.so``
(include/hbtgpu.h; sources in dynolog_b200/csrc/).  This package only holds
the ctypes plumbing used by tests and bench.py (``dynolog_b200.hbtgpu``) and
the synthetic capture generator (``dynolog_b200.synth``).
"""
__all__ = ["hbtgpu", "synth"]
