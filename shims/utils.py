"""`utils.clean_gpu_memory` (src/utils.py:5): collect garbage, drop torch's cached CUDA blocks, reset the
peak statistics, wait for the device. CUDA only -- this package has no MPS or CPU branch."""
import gc

import torch


def clean_gpu_memory():
    gc.collect()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
        torch.cuda.synchronize()
