"""fqtool_b200: B200-native per-read QC + trimming hot path of fqtool behind a C ABI.

Layout: csrc/ (CUDA kernels + the C ABI of include/fqk.h), abi.py (ctypes
mirror), options.py (host mirror of the reference's Options), batch.py (packed
read batches), engine.py (one context per GPU), processor.py (mirrors of
SingleEndProcessor / PairEndProcessor for one pack of reads).
"""
__version__ = "0.1.0"
