"""CPU suite: bag-data-parallel inference over variable-length bags (BASELINE config 5) -- the placement
is a partition, identical on every rank, balanced; the logits gather over gloo (world 2 and 3) returns every bag's row."""
import math
import os
import socket
import types

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from transmil_b200 import bagdp


def config5_lengths(n_bags=256, seed=5):
    """BASELINE config 5 / SURVEY 8d: bag lengths log-uniform in [1 000, 100 000], seed 5."""
    rng = np.random.default_rng(seed)
    return [int(v) for v in np.exp(rng.uniform(math.log(1000), math.log(100000), n_bags))]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("D", [1024, 2048])
def test_assignment_is_a_balanced_partition(world, D):
    lengths = config5_lengths()
    parts = bagdp.assign_bags(lengths, world, D)
    assert sorted(i for p in parts for i in p) == list(range(len(lengths)))
    assert parts == bagdp.assign_bags(list(lengths), world, D)                    # deterministic
    load = [sum(bagdp.bag_cost(lengths[i], D) for i in p) for p in parts]
    ideal = sum(load) / world
    assert max(load) <= 1.02 * ideal                                              # 256 bags: LPT is within 2 % of perfect
    rr = [sum(bagdp.bag_cost(lengths[i], D) for i in range(r, len(lengths), world)) for r in range(world)]
    assert max(load) <= max(rr)                                                   # never worse than round-robin
    assert all(p == sorted(p) for p in parts)


def test_cost_model_follows_the_survey():
    assert bagdp.bag_cost(50000, 2048) == pytest.approx(11.69e6 * 50000 + 13.2e9)
    assert bagdp.bag_cost(50000, 1024) == pytest.approx(7.49e6 * 50000 + 13.2e9)
    assert bagdp.bag_cost(1000, 512) < bagdp.bag_cost(1000, 1024)
    assert bagdp.assign_bags([], 3) == [[], [], []]
    assert bagdp.assign_bags([5, 5, 5], 2) == [[0, 2], [1]]                       # ties: lower index, lower rank
    with pytest.raises(ValueError):
        bagdp.assign_bags([1], 0)


def _fake_logits(bag, n_classes):
    """Host stand-in for the forward: any deterministic function of the bag."""
    s = bag.double().sum()
    return torch.stack([s * (k + 1) / bag.shape[0] for k in range(n_classes)]).float().reshape(1, n_classes)


def _bag(i, n, D=16):
    return torch.rand(n, D, generator=torch.Generator().manual_seed(100 + i))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lengths, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        model = types.SimpleNamespace(n_classes=3, in_features=16)
        loaded = []

        def load(i):
            loaded.append(i)
            return _bag(i, lengths[i])

        runner = lambda bags: (_fake_logits(b, 3) for b in bags)
        rows = bagdp.BagParallelInference(model, runner=runner)(lengths, load)
        out[rank] = (rows, loaded)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_logits_gather_over_gloo(world):
    lengths = [int(v) for v in np.random.default_rng(3).integers(1, 400, 23)]
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), lengths, out), nprocs=world, join=True)
        res = dict(out)
    want = torch.cat([_fake_logits(_bag(i, n), 3) for i, n in enumerate(lengths)])
    parts = bagdp.assign_bags(lengths, world, 16)
    for r in range(world):
        rows, loaded = res[r]
        assert torch.equal(rows, want)                      # every rank holds every row, exactly
        assert loaded == parts[r]                           # and read only its own bags


def test_single_rank_without_a_process_group():
    lengths = [7, 300, 12]
    model = types.SimpleNamespace(n_classes=2, in_features=16)
    rows = bagdp.BagParallelInference(model, runner=lambda bags: (_fake_logits(b, 2) for b in bags))(lengths, lambda i: _bag(i, lengths[i]))
    assert torch.equal(rows, torch.cat([_fake_logits(_bag(i, n), 2) for i, n in enumerate(lengths)]))
