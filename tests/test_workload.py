"""GPU: the device-generated synthetic workloads (bench.py's inputs) through the
device-pointer entry points, checked against the oracle on the unpacked ASCII."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cfg,n", [("cfg2", 60000), ("cfg4", 60000), ("cfg3", 20000), ("cfg5", 10000)])
def test_device_workload_matches_oracle(cfg, n):
    import torch
    from fqtool_b200 import abi, batch, workload
    from fqtool_b200.engine import Engine
    dev = torch.device("cuda", 0)
    opt = workload.options_for(cfg)
    eng = Engine(opt.to_params(), device=0)
    first = 123456
    b1, b2 = workload.generate(cfg, n, first, dev, eng.stream)
    o1 = torch.zeros(n * 16, dtype=torch.uint8, device=dev)
    o2 = torch.zeros(n * 16, dtype=torch.uint8, device=dev); op = torch.zeros(n * 24, dtype=torch.uint8, device=dev)
    s1 = b1.c_struct()
    if b2 is None:
        eng.process_se_device(s1, o1.data_ptr())
    else:
        eng.process_pe_device(s1, b2.c_struct(), o1.data_ptr(), o2.data_ptr(), op.data_ptr())
    eng.sync()
    flat = eng.finalize()
    dup = eng.dup_table() if opt.duplicate.enabled else None
    recs = [o1.cpu().numpy().view(abi.READ_RESULT_DTYPE)]
    if b2 is not None:
        recs += [o2.cpu().numpy().view(abi.READ_RESULT_DTYPE), op.cpu().numpy().view(abi.PAIR_RESULT_DTYPE)]
    h1 = b1.to_host(); r1 = batch.unpack(h1)
    r2 = batch.unpack(b2.to_host()) if b2 is not None else None
    # the generator emits exactly the layout the host packer produces
    p = batch.pack(r1, first)
    assert np.array_equal(p.qual, h1.qual) and np.array_equal(p.bases2, h1.bases2)
    assert np.array_equal(p.offsets, h1.offsets) and np.array_equal(p.lengths, h1.lengths)
    from oracle import pyoracle
    o = pyoracle.Oracle(opt.to_params(), "port")
    orec = (o.process_se(r1, first),) if r2 is None else o.process_pe(r1, r2, first)
    util.assert_same((tuple(recs), flat, dup), (orec, o.finalize(), o.dup_table() if opt.duplicate.enabled else None))
    # the workload exercises the path: some reads trimmed / failed
    assert (recs[0]["len"] != h1.lengths).any()
    eng.close()


def test_generator_is_counter_based():
    """Any chunk of the stream can be generated independently (sharding across ranks)."""
    import torch
    from fqtool_b200 import workload
    dev = torch.device("cuda", 0)
    a, _ = workload.generate("cfg2", 3000, 1000, dev, None)
    b, _ = workload.generate("cfg2", 1000, 2000, dev, None)
    torch.cuda.synchronize()
    Lp = 160
    assert torch.equal(a.qual[1000 * Lp:2000 * Lp], b.qual[:1000 * Lp])
    assert torch.equal(a.bases2[1000 * Lp // 4:2000 * Lp // 4], b.bases2[:1000 * Lp // 4])
