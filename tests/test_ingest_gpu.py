"""GPU suite: the slide loop with overlapped H2D copies returns exactly what direct calls return, in order, for bags
of different lengths (staging buffers are reused and regrown)."""
import pytest
import torch

from common import build_model, make_input

pytestmark = pytest.mark.gpu


def test_streamed_inference_matches_direct_calls(built_lib):
    import transmil_b200 as T
    m = build_model(3, 1024).cuda()
    bags = [make_input("rand", 1, n, 1024, 100 + i).pin_memory() for i, n in enumerate([3000, 700, 6000, 6000, 1, 2500, 9000])]
    with torch.no_grad():
        direct = [m(b.cuda()).cpu() for b in bags]
    loop = T.StreamedInference(m)
    got = list(loop(bags))
    assert len(got) == len(bags)
    for g, d in zip(got, direct):
        assert torch.equal(g, d)
    got2 = list(loop(b[0] for b in bags[:3]))          # 2-D bags, generator input, second use of the same loop object
    for g, d in zip(got2, direct[:3]):
        assert torch.equal(g, d)
    assert list(loop([])) == []
    with pytest.raises(ValueError):
        T.StreamedInference(m.train())
