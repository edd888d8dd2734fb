"""Drop-in proof: the C++ host adapter (fqtool_b200/host: packing, fqk_process_*, slicing, routing,
merged-read assembly) + the CUDA hot path must write byte-identical FASTQ files to the WHOLE
reference program (`fqtool_ref -w 1`, goldens in tests/golden/e2e_expected.json made by
tests/golden/make_e2e.py) for the same command lines."""
import hashlib
import json
import os
import subprocess

import pytest

from fqtool_b200 import build
from tests.golden import make_e2e

HERE = os.path.dirname(os.path.abspath(__file__))
EXPECTED = json.load(open(os.path.join(HERE, "golden", "e2e_expected.json")))
FQ1 = os.path.join(HERE, "golden", "fq", "r1.fq.gz")
FQ2 = os.path.join(HERE, "golden", "fq", "r2.fq.gz")


def test_harness_builds_and_rejects_out_of_scope_flags(tmp_path):
    exe = build.build_host()
    r = subprocess.run([exe, "-i", FQ1, "-o", str(tmp_path / "o.fq"), "--ora"], capture_output=True, text=True)
    assert r.returncode == 2 and "outside the hot-path scope" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(EXPECTED))
def test_output_files_match_reference_binary(case, tmp_path):
    exe = build.build_host()
    d = str(tmp_path)
    cmd = make_e2e.command(exe, case, FQ1, FQ2, d) + ["--max_item_in_pack", "700", "--max_cycles", "320"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    summary = json.loads(r.stdout)
    for name, want in EXPECTED[case]["files"].items():
        got = hashlib.md5(open(os.path.join(d, name), "rb").read()).hexdigest()
        assert got == want, f"{case}: {name} differs from the reference binary's output"
    fr = EXPECTED[case]["filter_result"]
    assert summary["filter"]["pass"] == fr["PassedFilterReads"]
    assert summary["filter"]["quality"] == fr["LowQualityReads"] and summary["filter"]["n_base"] == fr["TooManyNReads"]
