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


def _worker_round2(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sat_b200 import sharded
        from sat_b200.decoder import Decoder
        from sat_b200.trainer import FusedTrainStep
        # 1. replicas built from DIFFERENT seeds leave FusedTrainStep.__init__ with rank 0's weights (parameters were
        #    never broadcast, a seed skew drifted silently), frozen / unused tensors included
        torch.manual_seed(100 + rank)
        dec = Decoder(40, 64, tf=True, ado=True, bert=False, attention=True)
        before = dec.lstm.weight_ih.detach().clone()
        tr = FusedTrainStep(dec)
        flat = [torch.zeros_like(tr.flat_param) for _ in range(world)]
        dist.all_gather(flat, tr.flat_param)
        assert all(torch.equal(f, flat[0]) for f in flat)
        if rank != 0:
            assert not torch.equal(before, dec.lstm.weight_ih)
        unused = [torch.zeros_like(dec.deep_output.weight) for _ in range(world)]      # ado: deep_output is outside the flat buffer (Q7)
        dist.all_gather(unused, dec.deep_output.weight.data)
        assert all(torch.equal(u, unused[0]) for u in unused)
        assert abs(tr.hyper[5].item() - 1.0 / world) < 1e-9                            # the mean's 1/world rides in the Adam kernel
        # a mismatching layout is refused on every rank
        dec2 = Decoder(40, 64, tf=True, ado=(rank == 0), bert=False, attention=True)
        try:
            FusedTrainStep(dec2)
            raise AssertionError("layout mismatch not detected")
        except RuntimeError as e:
            assert "layout differs" in str(e)
        # 1b. bf16 gradient exchange: the sum arrives in the bf16 image Adam consumes; small integers are exact in bf16
        torch.manual_seed(7)
        tr16 = FusedTrainStep(Decoder(40, 64, tf=True, bert=False, attention=True), grad_exchange_dtype=torch.bfloat16)
        tr16.flat_grad.copy_((torch.arange(tr16.n) % 61).float() + rank)
        tr16.sync_grads()
        want = world * (torch.arange(tr16.n) % 61).float() + sum(range(world))
        assert tr16.reduced_grad().dtype == torch.bfloat16 and torch.equal(tr16.reduced_grad().float(), want)
        # 2. sharded beam search: unequal shards, padded gather on rank 0, image order = rank order
        lo, hi = sharded.shard_range(7, rank, world)
        assert (lo, hi) == ((0, 4) if rank == 0 else (4, 7))
        n, maxlen, Lloc, beam = hi - lo, 6, 5, 3
        raw = {"lens": torch.arange(lo, hi, dtype=torch.int32) % 4 + 2,
               "status": torch.tensor([1 if (i % 3 == 2) else 0 for i in range(lo, hi)], dtype=torch.int32),
               "live": torch.full((n,), 2, dtype=torch.int32),
               "tokens": (torch.arange(lo, hi).view(n, 1) * 10 + torch.arange(maxlen).view(1, maxlen)).to(torch.int64),
               "alphas": torch.arange(lo, hi, dtype=torch.float32).view(n, 1, 1).expand(n, maxlen, Lloc).contiguous(),
               "last_alpha": torch.arange(lo, hi, dtype=torch.float32).view(n, 1, 1).expand(n, beam, Lloc).contiguous() + 0.5}
        got = sharded.gather_caption_results(raw, dst=0)
        if rank == 0:
            res = Decoder.results_from_raw(got)
            assert len(res) == 7
            for i, (sent, alpha) in enumerate(res):
                if i % 3 == 2:                       # fallback (decoder.py:256-258): ([0], Tensor(live, L))
                    assert sent == [0] and torch.is_tensor(alpha) and alpha.shape == (2, Lloc) and alpha[0, 0].item() == i + 0.5
                else:
                    assert sent == [i * 10 + j for j in range(i % 4 + 2)]
                    assert len(alpha) == len(sent) and alpha[0][0] == float(i)
            tok_only = sharded.gather_caption_results(raw, dst=0, with_alphas=False)
            assert "alphas" not in tok_only and tok_only["tokens"].shape == (7, maxlen)
        else:
            assert got is None
            assert sharded.gather_caption_results(raw, dst=0, with_alphas=False) is None
        q.put((rank, "ok"))
    except Exception as e:                            # pragma: no cover - reported to the parent
        import traceback
        q.put((rank, traceback.format_exc() or repr(e)))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_replica_sync_and_sharded_caption_gather():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_round2, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
