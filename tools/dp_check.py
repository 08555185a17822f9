"""Data-parallel check on real GPUs (torchrun, >= 2 ranks): three fused train steps from identical weights with the gradient
all-reduce (a) as ONE call after the backward pass and (b) as two buckets, the early one overlapped with the backward tail
(FusedTrainStep.overlap_allreduce).  Parameters must agree across the two modes (same sums; NCCL may order a ring differently for
a different buffer split, so fp32 rounding only) and be identical on every rank.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_check.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, world, lr_ = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr_)
    dev = torch.device("cuda", lr_)
    dist.init_process_group("nccl", device_id=dev)
    from sat_b200 import synth
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D, B, L, T = 1003, 2048, 64, 196, 6
    P = synth.make_params(V, D, 512, seed=9, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=100 + rank)            # every rank its own shard
    a, cap = a.to(dev).to(torch.bfloat16), cap.to(dev)
    res = {}
    for overlap in (False, True):
        dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.bfloat16)
        dec.load_state_dict(P)
        dec = dec.to(dev).eval()                                          # no dropout: both modes see the same step
        tr = FusedTrainStep(dec, lr=1e-3)
        tr.overlap_allreduce = overlap
        for _ in range(3):
            loss = tr.step(a, cap)
        torch.cuda.synchronize()
        res[overlap] = (tr.flat_param.clone(), loss.clone())
        others = [torch.empty_like(tr.flat_param) for _ in range(world)]
        dist.all_gather(others, tr.flat_param)
        assert all(torch.equal(o, others[0]) for o in others), f"overlap={overlap}: parameters differ between ranks"
    # bf16 gradient exchange + CUDA-graph replay (what bench.py runs at N > 1): replicas stay bit-identical, and the parameters
    # stay within the update size of the fp32 exchange
    dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.bfloat16)
    dec.load_state_dict(P)
    dec = dec.to(dev).eval()
    tr = FusedTrainStep(dec, lr=1e-3, grad_exchange_dtype=torch.bfloat16, use_cuda_graph=True)
    for _ in range(5):
        loss16 = tr.step(a, cap)
    torch.cuda.synchronize()
    others = [torch.empty_like(tr.flat_param) for _ in range(world)]
    dist.all_gather(others, tr.flat_param)
    assert all(torch.equal(o, others[0]) for o in others), "bf16 exchange: parameters differ between ranks"
    if rank == 0:
        print(f"bf16 exchange + graph: {tr.graph_replays} graph replays, replicas identical, loss after 5 steps {loss16.tolist()}")
    assert tr.graph_replays == 3
    p0, p1 = res[False][0], res[True][0]
    err = (p0 - p1).abs().max().item()
    upd = 3 * 1e-3                                                        # Adam moves an element by <= ~lr per step
    if rank == 0:
        print(f"ranks={world} n_early={tr.n_early} of {tr.n} floats; max |param(one call) - param(two buckets)| = {err:.3e} "
              f"(three Adam steps of lr 1e-3); losses {res[False][1].tolist()} / {res[True][1].tolist()}")
    assert err <= 0.02 * upd, err
    # a live CUDA graph that captured NCCL work keeps the communicator busy: destroy_process_group() then hangs.  Drop the graphs first.
    tr.close()
    del tr
    torch.cuda.synchronize()
    dist.destroy_process_group()
    if rank == 0:
        print("dp_check OK")


if __name__ == "__main__":
    try:
        main()
    except BaseException:                      # a rank that dies while its peer sits in a collective would hang the job
        import traceback
        traceback.print_exc()
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(1)
