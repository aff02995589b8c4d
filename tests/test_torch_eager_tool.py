"""CPU suite: the stock-PyTorch baseline module of tools/bench_torch_eager.py is the same function as the
pinned CPU oracle, so that the GPU number it produces is a baseline for the same work."""
import importlib.util
import os

import pytest
import torch

from common import ROOT, build_model, make_input
from oracle import transmil_oracle as O


def _tool():
    spec = importlib.util.spec_from_file_location("_eager_tool", os.path.join(ROOT, "tools", "bench_torch_eager.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("D,nc,n,B", [(2048, 2, 300, 2), (1024, 3, 1000, 1), (768, 2, 257, 1), (512, 6, 5, 1)])
def test_stock_module_equals_oracle(D, nc, n, B):
    torch.set_float32_matmul_precision("highest")
    m = build_model(nc, D)
    stock = _tool().StockTransMIL(m).eval()
    assert list(stock.state_dict()) == list(m.state_dict())
    x = make_input("randn", B, n, D, 9)
    with torch.no_grad():
        got = stock(x)
    ref = O.forward(m.state_dict(), x)
    assert (got - ref).abs().max() <= 2e-6
