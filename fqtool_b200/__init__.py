"""fqtool_b200: B200-native per-read QC + trimming hot path of fqtool behind a C ABI.

Layout: csrc/ (CUDA kernels + the C ABI of include/fqk.h), abi.py (ctypes mirror), options.py (host mirror of the
reference's Options), batch.py (packed read batches), engine.py (one context per GPU: packed, ASCII, text and raw-FASTQ
entry points), prepass.py (adapter auto-detection: device counts + the host's key ranking / tree walk), report.py
(Stats::summarize / Duplicate::statAll on the raw arrays), workload.py + synth.py (synthetic BASELINE workloads),
distributed.py (NumPy mirror of the cross-rank merge for the gloo test), host/ (C++ adapter fqk::HotPath + harness).
The reference-side binding that is compiled into the reference program lives in integration/ at the repository root.
"""
__version__ = "0.2.0"
