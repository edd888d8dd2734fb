"""CPU checks of the C-ABI library: it loads, exports every symbol include/fqk.h declares,
its host-side pieces (packer, layout, defaults) behave, and it refuses to run without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from fqtool_b200 import abi, batch, synth
from fqtool_b200.options import Options, from_cli
from oracle import pyoracle
from tests import configs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    lib = abi.load()
    hdr = open(os.path.join(ROOT, "include", "fqk.h")).read()
    declared = set(re.findall(r"\b(fqk_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(abi.EXPORTS), declared ^ set(abi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_sizes_match_c(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "fqk.h"\nint main(){printf("%zu %zu %zu %zu %zu", sizeof(fq_params),'
                   'sizeof(fq_batch), sizeof(fq_read_result), sizeof(fq_pair_result), sizeof(fq_stats_layout));return 0;}')
    exe = tmp_path / "sz"
    assert os.system(f"gcc -I{ROOT}/include {src} -o {exe}") == 0
    got = [int(x) for x in os.popen(str(exe)).read().split()]
    assert got == [C.sizeof(abi.FqParams), C.sizeof(abi.FqBatch), abi.READ_RESULT_DTYPE.itemsize,
                   abi.PAIR_RESULT_DTYPE.itemsize, C.sizeof(abi.FqStatsLayout)]


def test_defaults_match_options_mirror():
    lib = abi.load()
    p = abi.FqParams()
    lib.fqk_default_params(C.byref(p))
    q = Options().to_params()
    for name, _ in abi.FqParams._fields_:
        assert getattr(p, name) == getattr(q, name), name
    assert q.low_quality_base_limit == 22 and q.low_quality_limit == 53     # options.cpp:26,44


def test_cli_quirks():
    o = from_cli("-W 8 -M 30 --enable_cut_front".split())                   # dead flags, SURVEY.md section 5
    assert o.qualitycut.windowSizeFront == 4 and o.qualitycut.qualityFront == 20
    assert from_cli([]).to_params().qual_filter_enabled == 0                 # CLI11 flag reset
    with pytest.raises(ValueError):
        from_cli(["--ora"])


@pytest.mark.parametrize("name", ["cfg2", "cfg4", "cfg3", "pe_dup_kmer"])
def test_layout_matches_oracle(name):
    lib = abi.load()
    p = configs.options(name).to_params()
    L = abi.FqStatsLayout()
    assert lib.fqk_stats_layout(C.byref(p), C.byref(L)) == 0
    Lo = pyoracle.layout(p)
    assert bytes(L) == bytes(Lo)


def test_pack_roundtrip_and_alignment():
    for reads in (synth.ragged(500, seed=5), synth.se_hiseq(100, 150, seed=6), batch.from_strings([], [])):
        b = batch.pack(reads)
        assert np.all(b.offsets % abi.FQK_PAD == 0) and b.total_bases % abi.FQK_PAD == 0
        s, q, off = batch.unpack(b)
        assert np.array_equal(s, reads[0]) and np.array_equal(q, reads[1]) and np.array_equal(off, reads[2])
        # N is stored as code 0 with the flag in quality bit 7
        nidx = np.nonzero(reads[0] == ord("N"))[0]
        assert int(((b.qual & 0x80) != 0).sum()) == nidx.size


def test_pack_rejects_foreign_alphabet():
    with pytest.raises(ValueError):
        batch.pack(batch.from_strings(["ACGTa"], ["IIIII"]))
    with pytest.raises(ValueError):
        batch.pack(batch.from_strings(["ACGTR"], ["IIIII"]))
    with pytest.raises(ValueError):
        batch.pack(batch.from_strings(["A" * 5000], ["I" * 5000]))


def test_no_cpu_fallback():
    """Without a GPU the product path must fail loudly (never route through the oracle)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fqtool_b200.engine import Engine, FqkError
    with pytest.raises(FqkError, match="no CUDA device"):
        Engine(Options().to_params())


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "fqtool_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in txt and "fq_oracle" not in txt and "libfqref" not in txt, f


def test_invalid_params_rejected():
    lib = abi.load()
    h = C.c_void_p()
    for mut in (dict(kmer_len=16), dict(dup_enabled=1, dup_keylen=20),
                dict(cut_front_enabled=1, cut_front_window=0), dict(abi_version=99)):
        p = Options().to_params()
        for k, v in mut.items():
            setattr(p, k, v)
        assert lib.fqk_create(C.byref(h), C.byref(p), 0, 0, 1, None) == -1, mut


def test_shipped_library_has_no_work_skipping_switches():
    """The timing / A-B switches of the kernels (FQK_V3_DEBUG, FQK_V4_DEBUG ...: some skip work) are compiled in only with
    -DFQK_TUNING_SWITCHES; the library the tests and bench.py load must not even contain their names."""
    from fqtool_b200 import build
    if "FQK_TUNING_SWITCHES" in os.environ.get("FQK_NVCC_EXTRA", ""):
        pytest.skip("tuning build")
    blob = open(build.LIB, "rb").read()
    for name in (b"FQK_V3_DEBUG", b"FQK_V4_DEBUG", b"FQK_V3_WIDE_FAST", b"FQK_V4_TMODE", b"FQK_KMER_FUSED", b"FQK_V4_SPLIT"):
        assert name not in blob, name
