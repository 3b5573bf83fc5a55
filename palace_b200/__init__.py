"""palace_b200 — B200-native implementation of PALACE's `eref` k-mer reference-identification step.

Only what that one path needs: csrc/ (sm_100a CUDA kernels + the C ABI of include/eref_gpu.h + the
drop-in `eref` executable), eref.py (host-side mirror of the reference's process interface over the
C ABI) and synth.py (seeded synthetic inputs).  No CPU fallback anywhere in this package.
"""
from .eref import (ErefError, ErefGPU, RESULT_DTYPE, format_results, load_library, random_coder,  # noqa: F401
                   read_coder, run_eref)
