"""fqk_prepass_kmer_counts_static (include/fqk.h): the pre-pass k-mer count for callers that must not allocate.

fqk_prepass_kmer_counts gets its working histogram from cudaMalloc on every call.  A thread that runs the pre-pass while
other threads have packs in flight on the same GPU cannot use that (cudaMalloc / cudaFree hold up every stream of the
device), and neither can a process whose HBM is already fully handed out (the allocation fails).  The static variant
keeps the histogram in a 4 MiB array built into libfqk.so and takes the reads and the output as device buffers of the
caller, so it allocates nothing.

What is checked, through the C ABI: its counts are bit-identical to fqk_prepass_kmer_counts' on the same reads (integer
counts: the tolerance of the other GPU tests is equality), on two calls in a row with different reads AND different
key lengths -- the array is shared by all calls and keeps the previous call's bins, so the second call would show stale
counts if the entry point did not clear what it uses; and a key too long for the array is refused with a message.

This file sorts last on purpose (pytest -x stops at the first failure; a failure here must not hide other tests).  It
needs no state from them: it passes alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

from fqtool_b200 import abi, prepass, synth


def _static_counts(lib, torch, reads, keylen, first_pos, shift_tail):
    """Upload the reads with torch (the TEST may allocate; the entry point does not) and run the static variant."""
    seq, _, off = reads
    off = np.ascontiguousarray(off, dtype=np.uint64)
    n = len(off) - 1
    d_seq = torch.from_numpy(np.ascontiguousarray(seq[int(off[0]):int(off[n])]).view(np.uint8).copy()).cuda()
    if d_seq.numel() == 0:
        d_seq = torch.zeros(16, dtype=torch.uint8, device="cuda")
    d_off = torch.from_numpy((off - off[0]).astype(np.int64)).cuda()              # same bits as uint64
    d_cnt = torch.full((1 << (2 * keylen),), -1, dtype=torch.int64, device="cuda")    # poison: every bin must be written
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.fqk_prepass_kmer_counts_static(0, n, d_seq.data_ptr(), d_off.data_ptr(), keylen, first_pos, shift_tail,
                                            d_cnt.data_ptr(), stream)
    assert rc == 0, (rc, lib.fqk_last_error(None))
    torch.cuda.synchronize()
    return d_cnt.cpu().numpy().view(np.uint64)


@pytest.mark.gpu
def test_static_counts_match_allocating_entry_point_twice_in_a_row():
    import torch
    lib = abi.load()
    # call 1: the reference's own setting (10-mers from position 20) on reads with adapters, polyX tails and N bases
    a = synth.se_adapter_polyx(20000, 100, seed=21)
    want_a = prepass.kmer_counts(a, 10, 20, 1)
    got_a = _static_counts(lib, torch, a, 10, 20, 1)
    assert want_a.sum() > 0 and np.array_equal(got_a, want_a)
    # call 2, straight after: other reads (ragged lengths, some shorter than a window) and a SHORTER key, so the low bins
    # of the shared array still hold call 1's counts when it starts
    b = synth.ragged(5000, seed=22)
    want_b = prepass.kmer_counts(b, 6, 0, 0)
    got_b = _static_counts(lib, torch, b, 6, 0, 0)
    assert want_b.sum() > 0 and np.array_equal(got_b, want_b)
    # and the first input again: same answer as before
    assert np.array_equal(_static_counts(lib, torch, a, 10, 20, 1), want_a)
    # no reads: all bins zero
    empty = (np.zeros(0, dtype=np.uint8), None, np.zeros(1, dtype=np.uint64))
    assert not _static_counts(lib, torch, empty, 4, 0, 0).any()


@pytest.mark.gpu
def test_static_refuses_a_key_longer_than_the_array():
    import torch
    lib = abi.load()
    d = torch.zeros(64, dtype=torch.int64, device="cuda")
    rc = lib.fqk_prepass_kmer_counts_static(0, 1, d.data_ptr(), d.data_ptr(), 11, 0, 0, d.data_ptr(), None)
    assert rc == -6                                                               # FQK_E_LIMIT
    msg = lib.fqk_last_error(None).decode()
    assert "keylen 11" in msg and "fqk_prepass_kmer_counts" in msg
    assert lib.fqk_prepass_kmer_counts_static(0, 1, d.data_ptr(), d.data_ptr(), 0, 0, 0, d.data_ptr(), None) == -1
