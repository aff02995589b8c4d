"""GPU suite: slides read from feature-bag files (the reference's `features` / `coords` HDF5 layout, feature_dataloader.py:
252-254) into pinned memory and streamed through the slide loop give exactly the logits of direct calls on the same arrays."""
import numpy as np
import pytest
import torch

import h5_writer as W
from common import build_model

pytestmark = pytest.mark.gpu


def test_files_to_logits_through_the_slide_loop(built_lib, tmp_path):
    import transmil_b200 as T
    m = build_model(2, 1024).cuda()
    rng = np.random.default_rng(7)
    paths, arrays = [], []
    for i, (n, dt) in enumerate([(1500, np.float32), (300, np.float32), (4000, np.float16), (2200, np.float32)]):
        x = rng.random((n, 1024)).astype(dt)
        c = rng.integers(0, 100000, (n, 2)).astype(np.int64)
        paths.append(W.write_h5(tmp_path / f"slide{i}.h5", {"features": (x, dict(chunks=(1, 1024), maxshape=(None, 1024))),
                                                             "coords": c}))
        arrays.append(x)
    with torch.no_grad():
        direct = [m(torch.from_numpy(x).unsqueeze(0).cuda()).cpu() for x in arrays]
    got = list(T.StreamedInference(m)(T.h5bag.iter_bags(paths)))
    assert len(got) == len(direct)
    for g, d in zip(got, direct):
        assert torch.equal(g, d)
    bag, xy = T.h5bag.read_bag(paths[0])
    assert bag.is_pinned() and xy.shape == (1500, 2)
