"""HostFakeQuant (pinned host -> chunked H2D / kernel / D2H pipeline -> pinned host) returns exactly what
fake_quant_weight_dynamic returns for the same tensors on the device."""
import pytest
import torch

from lightcompress_b200 import quant as Q
from lightcompress_b200.host import HostFakeQuant

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


@pytest.mark.parametrize('gran,kw', [('per_group', {'group_size': 128}), ('per_channel', {})])
def test_host_pipeline_equals_device_call(gran, kw):
    g = torch.Generator().manual_seed(0)
    q = Q.IntegerQuantizer(4, False, gran, **kw)
    shapes = [(300, 1024), (64, 4096), (1000, 512), (7, 128)]
    hin = [(torch.randn(s, generator=g) * 0.02).to(torch.bfloat16).pin_memory() for s in shapes]
    hout = [torch.empty_like(h).pin_memory() for h in hin]
    pipe = HostFakeQuant(q, DEV, chunk_bytes=96 * 1024, depth=3)     # many chunks, ring reused several times
    for _ in range(2):
        pipe.begin()
        pipe.run(hin, hout)
        pipe.end()
    torch.cuda.synchronize()
    assert pipe.launches > 2 * len(shapes)
    for h, o in zip(hin, hout):
        ref = q.fake_quant_weight_dynamic(h.to(DEV)).cpu()
        assert torch.equal(o.view(torch.int16), ref.view(torch.int16))


def test_host_pipeline_rejects_unpinned_and_2d_units():
    q = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    pipe = HostFakeQuant(q, DEV)
    with pytest.raises(ValueError):
        pipe.run([torch.zeros(4, 128, dtype=torch.bfloat16)], [torch.zeros(4, 128, dtype=torch.bfloat16)])
    with pytest.raises(NotImplementedError):
        HostFakeQuant(Q.IntegerQuantizer(8, True, 'per_tensor'), DEV)
