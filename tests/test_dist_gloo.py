"""Host-side logic of the data-parallel train step on CPU: world_size-2 `gloo` processes.

The CUDA kernels cannot run here; what is covered is everything around them that differs for N > 1: the flat
parameter / gradient layout must be identical on every rank (it is what gets all-reduced), parameters the reference
leaves at grad=None (Q6, Q7) must be left out consistently, `sync_grads()` must sum over the ranks, and the mean
(1/world, folded into the Adam kernel) of per-shard mean gradients must equal the global-batch gradient.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, cfg, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sat_b200.decoder import Decoder
        from sat_b200.trainer import FusedTrainStep
        torch.manual_seed(1234)                       # same init on every rank, as after a broadcast / checkpoint load
        dec = Decoder(cfg["V"], cfg["D"], tf=True, ado=cfg["ado"], bert=False, attention=cfg["attention"])
        tr = FusedTrainStep(dec)
        assert tr.distributed and tr.world == world
        # 1. identical layout on every rank, parameters are views of the flat buffer with reference shapes
        sig = tr.layout_signature()
        sigs = [None] * world
        dist.all_gather_object(sigs, sig)
        assert all(s == sigs[0] for s in sigs)
        offs = [o for _, o, _ in sig]
        assert all(o % 4 == 0 for o in offs) and offs == sorted(offs)
        unused = dec.unused_parameter_keys()
        assert not (set(tr.train_keys) & set(unused))
        for k, p in dec.named_parameters():
            if k in tr.grad_views:
                assert tr.grad_views[k].shape == p.shape
        # 2. the exchange step sums the whole flat buffer
        tr.flat_grad.copy_(torch.arange(tr.n, dtype=torch.float32) % 97 + 1000.0 * rank)
        tr.sync_grads()
        expect = world * (torch.arange(tr.n, dtype=torch.float32) % 97) + 1000.0 * sum(range(world))
        assert torch.equal(tr.flat_grad, expect)
        # 3. mean over ranks of per-shard mean gradients == gradient of the global-batch mean loss (equal shards):
        #    checked on the oracle's loss with the reference layout of the batch split
        from oracle import sat_oracle as O
        from sat_b200 import synth
        P = {k: v.clone().requires_grad_(True) for k, v in synth.make_params(cfg["V"], cfg["D"], 512, seed=3).items()}
        Bg = 2 * world
        a, cap = synth.make_inputs(Bg, cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=5)

        def grads(sl):
            for v in P.values():
                v.grad = None
            preds, alphas = O.decoder_forward(P, a[sl], cap[sl], tf=True, ado=False, attention=True, dropout_masks=None)
            O.train_loss(preds, alphas, cap[sl], 1.0).backward()
            return torch.cat([P[k].grad.reshape(-1) for k in sorted(P) if P[k].grad is not None])

        per = Bg // world
        local = grads(slice(rank * per, (rank + 1) * per))
        dist.all_reduce(local)
        local /= world
        if rank == 0:
            glob = grads(slice(0, Bg))
            torch.testing.assert_close(local, glob, rtol=2e-4, atol=1e-6)
        q.put((rank, "ok"))
    except Exception as e:                            # pragma: no cover - reported to the parent
        import traceback
        q.put((rank, traceback.format_exc() or repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("attention,ado", [(True, False), (False, True)])
def test_two_rank_gloo_flat_gradient_exchange(attention, ado):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    cfg = dict(V=50, D=64, L=9, T=4, attention=attention, ado=ado)
    procs = [ctx.Process(target=_worker, args=(r, world, port, cfg, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
