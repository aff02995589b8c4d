"""torchrun worker of tests/test_sharded_gpu.py::test_sharded_forward_multi_gpu (and of tools/gpu_round scripts): the
token-sharded forward (peer-memory transport, eager and graph-captured, and the NCCL transport) on WORLD_SIZE GPUs against the unsharded forward on rank 0 and the CPU oracle."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import LOGIT_TOL, PROB_TOL, build_model, make_input  # noqa: E402


def main():
    from oracle import transmil_oracle as O
    from transmil_b200.sharded import TokenShardedTransMIL
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok, keep = True, []
    for n, D in [(6000, 1024), (20000, 2048), (50000, 1024)]:
        m = build_model(2, D).cuda()
        x = make_input("rand", 1, n, D, 5)
        sharded = TokenShardedTransMIL(m)                     # transport="peer": one C call per rank, NVLink peer-memory exchanges
        got = sharded(x).cpu()
        local_rows = sharded.plan(n).take(x, rank)           # the loader-side form: only this rank's rows
        got2 = sharded(local_rows, n=n).cpu()
        graphed = TokenShardedTransMIL(m, cuda_graph=True)   # eager first call, captured second, replayed third
        g1, g2, g3 = graphed(x).cpu(), graphed(x).cpu(), graphed(x * 1.0).cpu()
        g_other = graphed(make_input("rand", 1, n, D, 6)).cpu()
        via_nccl = TokenShardedTransMIL(m, transport="nccl")(x).cpu()   # Python-issued schedule over torch.distributed
        status = sharded.comm.status() | graphed.comm.status()
        with torch.no_grad():
            plain = m(x.cuda()).cpu()
        peer_ok = (status == 0 and torch.equal(g1, got) and torch.equal(g2, got) and torch.equal(g3, got) and not torch.equal(g_other, got)
                   and (via_nccl - got).abs().max().item() < 4e-4)     # both are inside LOGIT_TOL of the oracle; measured up to 1.4e-4 apart
        if not peer_ok:
            print(f"rank {rank} n={n}: status={status} graph==eager {torch.equal(g2, got)} replay {torch.equal(g3, got)} "
                  f"|nccl-peer|={(via_nccl - got).abs().max().item():.2e}", flush=True)
        ok = ok and peer_ok
        keep.append((sharded, graphed))                      # regions stay mapped until every rank is done (freed at exit)
        gathered = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather_object(gathered, got)
        if rank == 0:
            ref = O.forward({k: v.cpu() for k, v in m.state_dict().items()}, x)
            e_ref = (got - ref).abs().max().item()
            e_prob = (torch.softmax(got, 1) - torch.softmax(ref, 1)).abs().max().item()
            e_plain = (got - plain).abs().max().item()
            same = all(torch.equal(g, got) for g in gathered) and torch.equal(got, got2)
            print(f"n={n} D={D} world={world}: |sharded-oracle|={e_ref:.2e} |dprob|={e_prob:.2e} |sharded-unsharded|={e_plain:.2e} ranks_agree={same}", flush=True)
            ok = ok and e_ref < LOGIT_TOL and e_prob < PROB_TOL and e_plain < 2e-4 and same
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if rank == 0 and ok:
        print("SHARDED-OK", flush=True)
    sys.exit(0 if int(flag) == 1 else 1)


if __name__ == "__main__":
    main()
