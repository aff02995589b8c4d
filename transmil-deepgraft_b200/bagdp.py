"""Bag-data-parallel inference over variable-length bags (BASELINE config 5).

Bags are independent, so whole bags go to ranks and the data path has no collective; the only exchange is the final
gather of `[n_bags, n_classes]` logits.  What matters with bag lengths of 1 000 ... 100 000 patches is the assignment:
round-robin leaves ranks idle behind whoever drew the long slides, so bags are placed longest-first on the least loaded
rank under the survey's cost model `cost(n) = a(D) * n + c` (D1: 11.69 / 7.49 MFLOP per token for 2048-d / 1024-d
features + 13.2 GFLOP per bag for the two pseudo-inverses).  Every rank computes the same assignment from the list of
lengths alone -- no coordinator, no communication.  In the reference this role is played by Lightning's
DistributedSampler over the test set (code/train.py:184, one slide per step, model_interface.py:316-324).
"""
import torch

__all__ = ["bag_cost", "assign_bags", "BagParallelInference"]

PINV_FLOP_PER_BAG = 13.2e9
FLOP_PER_TOKEN = {2048: 11.69e6, 1024: 7.49e6}


def bag_cost(n, in_features=1024):
    """Algorithmic FLOPs of one bag of n patches (asymptotic form)."""
    a = FLOP_PER_TOKEN.get(int(in_features))
    if a is None:                       # single-Linear stem of another width: 7.49 MFLOP/token minus/plus the stem's share
        a = 7.49e6 + 2.0 * 512 * (int(in_features) - 1024)
    return a * int(n) + PINV_FLOP_PER_BAG


def assign_bags(lengths, world, in_features=1024):
    """Longest-processing-time-first placement.  -> list (one per rank) of bag indices in the order the rank runs them
    (ascending index, i.e. dataset order).  Deterministic: ties go to the lower bag index / lower rank."""
    if world < 1:
        raise ValueError("world must be >= 1")
    order = sorted(range(len(lengths)), key=lambda i: (-bag_cost(lengths[i], in_features), i))
    load = [0.0] * world
    mine = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += bag_cost(lengths[i], in_features)
        mine[r].append(i)
    return [sorted(m) for m in mine]


class BagParallelInference:
    """`BagParallelInference(model)(lengths, load)` -> logits `[n_bags, n_classes]` (host tensor, same on every rank).

    lengths: patches per bag, known up front (the reference's dataset knows them from the feature files' shapes);
    load(i): host bag i `[n_i, D]`, called only on the rank that owns bag i (e.g. `h5bag.read_bag(paths[i])[0]`).
    Each rank streams its own bags through `StreamedInference` (next bag's H2D under the current forward); the logits
    rows are gathered by one allreduce(sum) of a zero-filled `[n_bags, n_classes]` tensor -- exact, since exactly one
    rank writes each row.  Without an initialised process group it is a one-rank loop."""

    def __init__(self, model, group=None, runner=None):
        self.model, self.group = model, group
        self._runner = runner            # tests inject a host stand-in for the slide loop; product: StreamedInference

    def _world(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(self.group), dist.get_world_size(self.group)
        return 0, 1

    def __call__(self, lengths, load):
        rank, world = self._world()
        mine = assign_bags(lengths, world, getattr(self.model, "in_features", 1024))[rank]
        runner = self._runner
        if runner is None:
            from .ingest import StreamedInference
            runner = StreamedInference(self.model)
        n_classes = int(self.model.n_classes)
        device = self.model.cls_token.device if self._runner is None else torch.device("cpu")
        rows = torch.zeros((len(lengths), n_classes), dtype=torch.float32)
        for i, logits in zip(mine, runner(load(i) for i in mine)):
            rows[i] = logits.reshape(n_classes).float()
        if world > 1:
            import torch.distributed as dist
            buf = rows.to(device)         # NCCL reduces device tensors; gloo (CPU tests) host tensors
            dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
            rows = buf.cpu()
        return rows
