"""B200-native (sm_100a) intervention-and-readout path of lmmontoya-ai/taboo-brittleness.

Host side: Python + PyTorch tensors behind the reference's own surface (``models``,
``feature_map``, ``metrics``); arithmetic: hand-written CUDA kernels in ``csrc/`` behind the
C ABI of ``include/tb_b200.h`` (``libtb_b200.so``, bound with ctypes in ``_lib``/``ops``).
There is no CPU fallback.
"""
__version__ = "0.1.0"
