"""Device-side output compaction and FASTQ record assembly (fqk_process_*_text): the text blocks
per output channel, concatenated over packs, must be byte-identical to the files the WHOLE reference program writes
(`fqtool_ref -w 1`, md5 in tests/golden/e2e_expected.json made by tests/golden/make_e2e.py) for the same command lines:
plain, failed (tagged), unpaired and merged outputs, base correction included.  No per-read string work on the host."""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

from fqtool_b200 import abi
from fqtool_b200.options import from_cli
from tests.golden import make_e2e

HERE = os.path.dirname(os.path.abspath(__file__))
EXPECTED = json.load(open(os.path.join(HERE, "golden", "e2e_expected.json")))
FQ = [os.path.join(HERE, "golden", "fq", f"r{i}.fq.gz") for i in (1, 2)]
FILES = {"out1.fq": abi.CH_OUT1, "out2.fq": abi.CH_OUT2, "merged.fq": abi.CH_MERGED, "failed.fq": abi.CH_FAILED,
         "un1.fq": abi.CH_UNPAIRED1, "un2.fq": abi.CH_UNPAIRED2}


def parse_fastq(path):
    """What FqReader holds (src/fqreader.cpp:90-195): four lines per record, kept as flat buffers + offsets."""
    lines = gzip.open(path, "rb").read().split(b"\n")
    recs = [lines[i:i + 4] for i in range(0, len(lines) - 3, 4)]
    return recs


def flat(recs, lo, hi):
    def cat(k):
        parts = [r[k] for r in recs[lo:hi]]
        off = np.zeros(len(parts) + 1, dtype=np.uint64)
        off[1:] = np.cumsum([len(p) for p in parts])
        return np.frombuffer(b"".join(parts), dtype=np.uint8), off
    names, noff = cat(0); seq, off = cat(1); strands, soff = cat(2); qual, _ = cat(3)
    return names, noff, seq, qual, off, strands, soff


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(EXPECTED))
def test_text_blocks_match_reference_files(case):
    from fqtool_b200.engine import Engine
    paired, flags = make_e2e.CASES[case]
    o = from_cli(flags.split() + (["-I", "r2"] if paired else []))
    o.maxCycles = 320
    eng = Engine(o.to_params())
    r1 = parse_fastq(FQ[0]); r2 = parse_fastq(FQ[1]) if paired else None
    outs = make_e2e.outputs(case, paired, "")
    writers = (abi.W_FAILED if "--failed_out" in outs else 0) | (abi.W_UNPAIRED1 if "--unpaired_read1" in outs else 0)
    acc = [b""] * abi.CH_COUNT
    nrec_total = [0] * abi.CH_COUNT
    for lo in range(0, len(r1), 700):
        hi = min(len(r1), lo + 700)
        if paired:
            blocks, nrec, passed, _ = eng.process_pe_text(flat(r1, lo, hi), flat(r2, lo, hi), writers, lo)
        else:
            blocks, nrec, passed, _ = eng.process_se_text(flat(r1, lo, hi), writers, lo)
        for ch in range(abi.CH_COUNT):
            acc[ch] += blocks[ch]; nrec_total[ch] += nrec[ch]
    eng.close()
    for name, want in EXPECTED[case]["files"].items():
        ch = FILES[name]
        assert hashlib.md5(acc[ch]).hexdigest() == want, f"{case}: {name} differs from the reference binary's output"
        assert acc[ch].count(b"\n") == 4 * nrec_total[ch]


@pytest.mark.gpu
def test_text_buffer_too_small_reports_needed_size():
    import ctypes as C
    from fqtool_b200.engine import Engine
    o = from_cli("-q".split()); o.maxCycles = 160
    eng = Engine(o.to_params())
    r1 = parse_fastq(FQ[0])
    b, keep = eng._text(flat(r1, 0, 500), 0)
    outs = (abi.FqTextOut * abi.CH_COUNT)()
    small = np.zeros(64, dtype=np.uint8)
    outs[abi.CH_OUT1].data = abi.ptr(small); outs[abi.CH_OUT1].cap = 64
    rc = eng.lib.fqk_process_se_text(eng.h, C.byref(b), 0, None, outs, None)
    assert rc == -6 and outs[abi.CH_OUT1].len > 64 and outs[abi.CH_OUT1].records > 0
    eng.close()


# ---------------------------------------------------------------- raw FASTQ blocks
def _stream_fastq(eng, paired, texts, writers, block):
    """Feed decompressed text in blocks cut at ARBITRARY byte positions; the unconsumed tail is carried over."""
    acc = [b""] * abi.CH_COUNT
    pos = [0, 0]; first = 0
    while True:
        chunk = [t[p:p + block] for t, p in zip(texts, pos)]
        if paired:
            blocks, n, used, _ = eng.process_pe_fastq(chunk[0], chunk[1], writers, first)
        else:
            blocks, n, u, _ = eng.process_se_fastq(chunk[0], writers, first)
            used = (u,)
        for ch in range(abi.CH_COUNT):
            acc[ch] += blocks[ch]
        first += n
        for k, u in enumerate(used):
            pos[k] += u
        if n == 0:
            assert all(len(c) < block for c in chunk), "a block without one complete record: enlarge the block"
            break
    return acc, first


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(EXPECTED))
def test_raw_fastq_blocks_match_reference_files(case):
    """Decompressed FASTQ text in, FASTQ text out: newline scan, packing, hot path and record assembly on the device."""
    from fqtool_b200.engine import Engine
    paired, flags = make_e2e.CASES[case]
    o = from_cli(flags.split() + (["-I", "r2"] if paired else []))
    o.maxCycles = 320
    eng = Engine(o.to_params())
    texts = [gzip.open(f, "rb").read() for f in (FQ if paired else FQ[:1])]
    outs = make_e2e.outputs(case, paired, "")
    writers = (abi.W_FAILED if "--failed_out" in outs else 0) | (abi.W_UNPAIRED1 if "--unpaired_read1" in outs else 0)
    acc, n = _stream_fastq(eng, paired, texts, writers, 250_007)
    eng.close()
    assert n == 2000
    for name, want in EXPECTED[case]["files"].items():
        assert hashlib.md5(acc[FILES[name]]).hexdigest() == want, f"{case}: {name} differs from the reference binary's output"


@pytest.mark.gpu
def test_raw_fastq_crlf_and_malformed():
    from fqtool_b200.engine import Engine, FqkError
    o = from_cli("-q --enable_cut_tail".split()); o.maxCycles = 160
    text = gzip.open(FQ[0], "rb").read()
    eng = Engine(o.to_params())
    want, n, used, passed = eng.process_se_fastq(text)
    assert n == 2000 and used == len(text) and passed == want[abi.CH_OUT1].count(b"\n") // 4
    got, n2, used2, _ = eng.process_se_fastq(text.replace(b"\n", b"\r\n"))      # FqReader::getLine drops the '\r' too
    assert n2 == 2000 and got[abi.CH_OUT1] == want[abi.CH_OUT1]
    part, n3, used3, _ = eng.process_se_fastq(text[:100_000])                    # ends inside a record
    assert 0 < n3 < 2000 and text[used3 - 1:used3] == b"\n" and want[abi.CH_OUT1].startswith(part[abi.CH_OUT1])
    assert eng.process_se_fastq(b"")[1] == 0
    eng.close()
    lines = text.split(b"\n")
    lines[6] = b"-"                                                              # third line of the second record
    eng = Engine(o.to_params())
    with pytest.raises(FqkError):
        eng.process_se_fastq(b"\n".join(lines))
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["cfg2_r1", "cfg4_r1", "se_failed"])
def test_pipelined_text_matches_reference_files(case):
    """fqk_submit_se_text / fqk_wait_text with two packs in flight: same files as the reference program, same accumulators
    as the synchronous call."""
    from fqtool_b200.engine import Engine
    paired, flags = make_e2e.CASES[case]
    o = from_cli(flags.split()); o.maxCycles = 320
    r1 = parse_fastq(FQ[0])
    outs = make_e2e.outputs(case, paired, "")
    writers = abi.W_FAILED if "--failed_out" in outs else 0
    eng = Engine(o.to_params())
    acc = [b""] * abi.CH_COUNT
    pending = None
    for lo in list(range(0, len(r1), 450)) + [None]:
        h = eng.submit_se_text(flat(r1, lo, min(len(r1), lo + 450)), writers, lo) if lo is not None else None
        if pending is not None:
            blocks, nrec, passed, recs = eng.wait_text(pending)
            assert passed == nrec[abi.CH_OUT1]
            for ch in range(abi.CH_COUNT):
                acc[ch] += blocks[ch]
        pending = h
    flat_pipe = eng.finalize()
    eng.close()
    for name, want in EXPECTED[case]["files"].items():
        assert hashlib.md5(acc[FILES[name]]).hexdigest() == want, f"{case}: {name} differs from the reference binary's output"
    eng = Engine(o.to_params())
    for lo in range(0, len(r1), 450):
        eng.process_se_text(flat(r1, lo, min(len(r1), lo + 450)), writers, lo)
    assert np.array_equal(eng.finalize(), flat_pipe)
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["cfg2_r1", "se_failed"])
def test_pipelined_fastq_blocks_match_reference_files(case):
    """fqk_submit_se_fastq / fqk_wait_fastq, two blocks in flight; blocks cut at record boundaries by the caller."""
    from fqtool_b200.engine import Engine
    paired, flags = make_e2e.CASES[case]
    o = from_cli(flags.split()); o.maxCycles = 320
    text = gzip.open(FQ[0], "rb").read()
    lines = text.split(b"\n")[:-1]                                      # the text ends with a newline
    blocks_in = [b"\n".join(lines[i:i + 4 * 300]) + b"\n" for i in range(0, len(lines), 4 * 300)]
    assert b"".join(blocks_in) == text
    outs = make_e2e.outputs(case, paired, "")
    writers = abi.W_FAILED if "--failed_out" in outs else 0
    eng = Engine(o.to_params())
    acc = [b""] * abi.CH_COUNT
    pending, first, total = None, 0, 0
    for blk in blocks_in + [None]:
        h = eng.submit_se_fastq(blk, first) if blk is not None else None
        if blk is not None:
            first += 300
        if pending is not None:
            blocks, n, used, passed = eng.wait_fastq(pending, writers)
            assert used == len(pending["keep"][0]) and n > 0
            total += n
            for ch in range(abi.CH_COUNT):
                acc[ch] += blocks[ch]
        pending = h
    eng.close()
    assert total == 2000
    for name, want in EXPECTED[case]["files"].items():
        assert hashlib.md5(acc[FILES[name]]).hexdigest() == want, f"{case}: {name} differs from the reference binary's output"


@pytest.mark.gpu
def test_raw_fastq_truncated_quality_line_is_refused_cleanly():
    """A record whose quality line is shorter than its sequence (here: the last record of the block, so the sequence length
    would lead the packer past the uploaded bytes) is refused like the reference's reader does (src/fqreader.cpp:183-190),
    without a device fault: the same context and a new one keep working afterwards."""
    from fqtool_b200.engine import Engine, FqkError
    o = from_cli("-q --enable_cut_tail".split()); o.maxCycles = 160
    text = gzip.open(FQ[0], "rb").read()
    lines = text.split(b"\n")[:4 * 500]
    good = b"\n".join(lines) + b"\n"
    lines[-1] = lines[-1][:3]                                                    # 3 quality characters for 150 bases
    bad = b"\n".join(lines) + b"\n"
    eng = Engine(o.to_params())
    want = eng.process_se_fastq(good)
    with pytest.raises(FqkError):
        eng.process_se_fastq(bad)
    eng.close()
    eng = Engine(o.to_params())
    got = eng.process_se_fastq(good)
    eng.close()
    assert got[1] == want[1] == 500 and got[0][abi.CH_OUT1] == want[0][abi.CH_OUT1]


@pytest.mark.gpu
def test_merged_name_of_a_read_name_that_starts_with_a_space():
    """OverlapAnalysis::merge (src/overlapanalysis.cpp:95-100) splices the tag at name.substr(0, pos - 1); with the first
    space at position 0 that is substr(0, npos): the WHOLE name, then the tag, then the whole name again."""
    import re
    from fqtool_b200.engine import Engine
    o = from_cli("-m --discard_unmerged -q -I r2".split()); o.maxCycles = 320
    r1 = parse_fastq(FQ[0])[:600]; r2 = parse_fastq(FQ[1])[:600]
    eng = Engine(o.to_params())
    a, nrec_a, _, out_a = eng.process_pe_text(flat(r1, 0, 600), flat(r2, 0, 600), 0, 0)
    s1 = [[b" " + r[0]] + r[1:] for r in r1]
    b, nrec_b, _, out_b = eng.process_pe_text(flat(s1, 0, 600), flat(r2, 0, 600), 0, 0)
    eng.close()
    merged = [i for i in range(600) if out_a[2]["flags"][i] & 0x10]
    la = a[abi.CH_MERGED].split(b"\n"); lb = b[abi.CH_MERGED].split(b"\n")
    assert nrec_a[abi.CH_MERGED] == nrec_b[abi.CH_MERGED] == len(merged) > 50 and len(la) == len(lb) == 4 * len(merged) + 1
    for j, i in enumerate(merged):
        tag = re.search(rb"_merged_\d+_\d+", la[4 * j]).group(0)
        assert lb[4 * j] == s1[i][0] + tag + s1[i][0]
        assert lb[4 * j + 1:4 * j + 4] == la[4 * j + 1:4 * j + 4]
