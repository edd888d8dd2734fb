"""Golden vectors produced by the reference build (tests/golden/make_golden.py):
CPU: the oracle port reproduces them; GPU: the CUDA path through the C ABI reproduces them."""
import os

import pytest

from tests import configs, util

FILES = util.golden_files()
IDS = [os.path.basename(f)[:-4] for f in FILES]


def test_golden_present():
    assert len(FILES) >= 30


@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_oracle_port_matches_golden(path):
    g = util.load_golden(path)
    opt = configs.options(g["cfg"])
    util.assert_matches_golden(util.run_oracle(opt, g["r1"], g["r2"], chunks=2), g)


@pytest.mark.gpu
@pytest.mark.parametrize("impl", [0, 1, 3], ids=["auto", "generic", "fused_ws"])
@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_cuda_matches_golden(path, impl):
    g = util.load_golden(path)
    opt = configs.options(g["cfg"])
    util.assert_matches_golden(util.run_cuda(opt, g["r1"], g["r2"], chunks=2, impl=impl), g)
