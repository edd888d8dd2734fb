"""Known answers of the WHOLE reference binary (`fqtool_ref -w 1`) on its own testdata,
recorded at survey time (BASELINE.md section 5 / SURVEY.md section 8c).  They pin the oracle
against the reference program itself, independently of oracle/ref_driver.cpp.  CPU only."""
import os

import numpy as np
import pytest

from fqtool_b200 import batch, report
from oracle import pyoracle
from tests import configs

TD = "/root/reference/testdata"
pytestmark = pytest.mark.skipif(not os.path.exists(TD), reason="reference testdata absent")


def run(opt, r1, r2=None, kind="port"):
    o = pyoracle.Oracle(opt.to_params(), kind)
    rec = o.process_se(r1) if r2 is None else o.process_pe(r1, r2)
    rs = report.RunStats(o.finalize(), o.layout)
    dup = o.dup_table() if opt.duplicate.enabled else None
    return rec, rs, dup


@pytest.fixture(scope="module")
def td():
    return batch.read_fastq(f"{TD}/r1.fq.gz"), batch.read_fastq(f"{TD}/r2.fq.gz")


def test_cfg1(td):
    _, rs, _ = run(configs.options("cfg1"), *td)
    pre = [rs.stats[n].summarize() for n in ("pre1", "pre2")]
    post = [rs.stats[n].summarize() for n in ("post1", "post2")]
    assert sum(s["reads"] for s in pre) == 25000 and sum(s["bases"] for s in pre) == 3750000
    assert sum(s["q20"] for s in pre) == 3454793 and sum(s["q30"] for s in pre) == 3302399
    assert sum(s["reads"] for s in post) == 20026 and sum(s["bases"] for s in post) == 2931123
    assert sum(s["q20"] for s in post) == 2848927 and sum(s["q30"] for s in post) == 2762978
    assert rs.filter[0] == 20026 and rs.filter[20] == 4574 and rs.filter[12] == 24
    assert rs.polyx_reads[3] == 25000 and rs.polyx_bases[3] == 49014      # PE polyG argument-order quirk
    assert rs.adapter_reads == 1976 and rs.adapter_bases == 26412
    assert int(rs.insert_hist[:-1].sum()) == 10019 and int(rs.insert_hist[-1]) == 2481


def test_cfg2_flags_on_r1(td):
    _, rs, _ = run(configs.options("cfg2"), td[0])
    assert rs.filter[0] == 11284 and rs.filter[20] == 1174 and rs.filter[12] == 42
    assert rs.stats["post1"].summarize()["bases"] == 1691344


def test_cfg3_flags(td):
    _, rs, _ = run(configs.options("cfg3"), *td)
    assert rs.filter[0] == 24874
    post = [rs.stats[n].summarize() for n in ("post1", "post2")]
    assert sum(s["reads"] for s in post) == 15133 and sum(s["bases"] for s in post) == 2693477


def test_cfg4_flags_on_r1(td):
    _, rs, dup = run(configs.options("cfg4"), td[0])
    assert rs.filter[0] == 12499
    assert list(rs.polyx_reads) == [4, 4, 7, 12, 54] and list(rs.polyx_bases) == [69, 41, 82, 822, 1450]
    assert rs.adapter_reads == 928
    hist, _, rate = report.dup_stat_all(dup[1], dup[2], 32)
    assert list(hist[1:6]) == [5086, 150, 35, 13, 2]
    assert rate == 0.051130247578040904
