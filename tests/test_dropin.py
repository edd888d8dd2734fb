"""The literal drop-in: the reference's own program with the bodies of
SingleEndProcessor::processSingleEnd / PairEndProcessor::processPairEnd replaced by calls into libfqk.so
(integration/, built as oracle/_ref/fqtool_fqk by `make -C oracle dropin`) against the unmodified reference program
(oracle/_ref/fqtool_ref -w 1) on the reference's own test data: every output file byte-identical and report.json
identical except for the "Software" block (command line and working directory).  That compares, through
JsonReporter (src/jsonreporter.cpp:23-162), everything the hot path feeds: summary counts, filter result counters,
adapter-text maps, polyX tallies, duplication rate/histogram/mean GC, insert-size histogram and peak, the per-cycle
quality/content curves and the k-mer table of all four Stats objects."""
import json
import os
import subprocess

import pytest

from tests.golden import make_e2e

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(ROOT, "oracle", "_ref", "fqtool_ref")
FQK = os.path.join(ROOT, "oracle", "_ref", "fqtool_fqk")
FQ1 = os.path.join(HERE, "golden", "fq", "r1.fq.gz")
FQ2 = os.path.join(HERE, "golden", "fq", "r2.fq.gz")

EXTRA = {
    # option sets beyond tests/golden/make_e2e.py: duplication + k-mers on pairs, explicit adapters + polyX on pairs
    "pe_dup_kmer": (True, "-d --kmer --kmer_length 5 -q -F 2 -T 3 -f 1"),
    "pe_adapters_polyx": (True, f"-a --adapter_of_read1 {make_e2e.A1} --adapter_of_read2 {make_e2e.A2} -x -q --enable_cut_front --enable_cut_tail -B 120 -b 130"),
    "se_polyg_polyx": (False, "-g --min_len_detect_polyG 6 -x --base_to_trim AGN --min_len_detect_polyX 5 -q"),
    "cfg5": (True, "-q -g -a"),
}
CASES = dict(make_e2e.CASES, **EXTRA)


def _have():
    return os.path.exists(REF) and os.path.exists(FQK)


def test_seam_generator_anchors():
    """make_seam.py finds each anchor exactly once in the reference's sources (when they are present)."""
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("reference sources absent")
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        r = subprocess.run(["python", os.path.join(ROOT, "integration", "make_seam.py"), "/root/reference/src", d],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        for f in ("seprocessor_fqk.cpp", "peprocessor_fqk.cpp"):
            t = open(os.path.join(d, f)).read()
            assert "fqkseam::finish" in t and "_body.inc" in t
            assert "trimAndCut" not in t and "statRead" not in t.split("Stats::merge")[0].split("_body.inc")[-1]


def test_dropin_binary_links_against_libfqk():
    if not _have():
        pytest.skip("drop-in binaries not built (reference sources absent)")
    r = subprocess.run(["ldd", FQK], capture_output=True, text=True)
    assert "libfqk.so" in r.stdout and "not found" not in r.stdout


def _run(exe, case, d, pack, seam_text=None, workers=1):
    paired, flags = CASES[case]
    os.makedirs(d, exist_ok=True)
    env = dict(os.environ)
    if seam_text is not None:
        env["FQK_SEAM_TEXT"] = str(seam_text)
    make_e2e.CASES.setdefault(case, CASES[case])
    cmd = make_e2e.command(exe, case, FQ1, FQ2, d) + ["-w", str(workers), "-J", os.path.join(d, "report.json"),
                                                       "-H", os.path.join(d, "report.html"), "--max_item_in_pack", str(pack)]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=d, env=env)
    assert r.returncode == 0, f"{' '.join(cmd)}\n{r.stdout}\n{r.stderr}"
    return r


@pytest.mark.gpu
@pytest.mark.parametrize("seam_text", [1, 0], ids=["device_text", "host_strings"])
@pytest.mark.parametrize("case", sorted(CASES))
def test_dropin_matches_reference_program(case, seam_text, tmp_path):
    """seam_text 1: the worker bodies get the output TEXT from the device (fqk_process_*_text, the default);
    0: they build one std::string per read on the host from the records (fqk_process_*_ascii)."""
    if not _have():
        pytest.skip("drop-in binaries not built (reference sources absent)")
    dr, dg = str(tmp_path / "ref"), str(tmp_path / "fqk")
    _run(REF, case, dr, 700)
    _run(FQK, case, dg, 700, seam_text)
    names = sorted(f for f in os.listdir(dr) if f.endswith(".fq"))
    assert names, "the reference wrote no output"
    for f in names:
        a, b = open(os.path.join(dr, f), "rb").read(), open(os.path.join(dg, f), "rb").read()
        assert a == b, f"{case}: {f} differs ({len(a)} vs {len(b)} bytes)"
    ja, jb = json.load(open(os.path.join(dr, "report.json"))), json.load(open(os.path.join(dg, "report.json")))
    ja.pop("Software"); jb.pop("Software")
    if not CASES[case][0] and "Duplication" in ja:
        # SingleEndProcessor::process zeroes dupHist with sizeof(int) * histSize (src/seprocessor.cpp:244-245): the upper
        # half of the size_t array is uninitialised heap in the reference itself, so only the lower half is comparable
        for j in (ja, jb):
            j["Duplication"]["Histogram"] = j["Duplication"]["Histogram"][:len(j["Duplication"]["Histogram"]) // 2]
    for k in ja:
        assert ja[k] == jb.get(k), f"{case}: report.json section {k} differs"
    assert set(ja) == set(jb)


@pytest.mark.gpu
def test_dropin_adapter_autodetection(tmp_path):
    """Evaluator::evaluateAdapterSeq needs 10000 records: the 2000 test pairs six times over.  The pre-pass of the drop-in
    (integration/evaluator_body.inc: 10-mer counts and seed hits from fqk_prepass_*) must report the same detected adapters
    (here: none for read 1, TruSeq universal for read 2) and the whole run must stay identical."""
    if not _have():
        pytest.skip("drop-in binaries not built (reference sources absent)")
    import gzip
    big = []
    for i, src in enumerate((FQ1, FQ2)):
        p = str(tmp_path / f"big{i + 1}.fq")
        data = gzip.open(src, "rb").read()
        open(p, "wb").write(data * 6)
        big.append(p)
    outs = {}
    for tag, exe in (("ref", REF), ("fqk", FQK)):
        d = str(tmp_path / tag); os.makedirs(d)
        cmd = [exe, "-i", big[0], "-I", big[1], "-o", f"{d}/out1.fq", "-O", f"{d}/out2.fq", "-w", "1", "-J", f"{d}/report.json",
               "-H", f"{d}/report.html", "-a", "-q", "-g"]
        r = subprocess.run(cmd, capture_output=True, text=True, cwd=d)
        assert r.returncode == 0, r.stderr
        j = json.load(open(f"{d}/report.json")); j.pop("Software")
        outs[tag] = (j, open(f"{d}/out1.fq", "rb").read(), open(f"{d}/out2.fq", "rb").read())
    assert outs["ref"][0]["AdapterTrim"]["Read2AdapterSequence"].startswith("AGATCGGAAGAGC")     # the pre-pass did find one
    assert outs["ref"][0]["AdapterTrim"] == outs["fqk"][0]["AdapterTrim"]
    assert outs["ref"] == outs["fqk"]


def _records(path):
    lines = open(path, "rb").read().split(b"\n")
    return sorted(tuple(lines[i:i + 4]) for i in range(0, len(lines) - 3, 4))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["cfg5", "pe_dup_kmer", "cfg4_r1"])
def test_dropin_with_four_workers_matches_order_independent_results(case, tmp_path):
    """-w 4: the reference's worker threads take the seam's mutex in whatever order they are scheduled, so packs reach the
    device out of input order.  Everything that is a sum over reads must still equal the
    reference's -w 1 run: every report.json section except Duplication.MeanGC (first-read-wins per key, so it follows the
    order) and the output files as SETS of records (the writers see the packs in the same scrambled order)."""
    if not _have():
        pytest.skip("drop-in binaries not built (reference sources absent)")
    dr, dg = str(tmp_path / "ref"), str(tmp_path / "fqk")
    _run(REF, case, dr, 150)
    _run(FQK, case, dg, 150, workers=4)
    names = sorted(f for f in os.listdir(dr) if f.endswith(".fq"))
    assert names
    for f in names:
        assert _records(os.path.join(dr, f)) == _records(os.path.join(dg, f)), f"{case}: {f} holds different records"
    ja, jb = json.load(open(os.path.join(dr, "report.json"))), json.load(open(os.path.join(dg, "report.json")))
    for j in (ja, jb):
        j.pop("Software")
        if "Duplication" in j:
            j["Duplication"].pop("MeanGC")
            if not CASES[case][0]:
                j["Duplication"]["Histogram"] = j["Duplication"]["Histogram"][:len(j["Duplication"]["Histogram"]) // 2]
    assert set(ja) == set(jb)
    for k in ja:
        assert ja[k] == jb[k], f"{case}: report.json section {k} differs"
