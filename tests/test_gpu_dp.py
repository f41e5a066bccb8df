"""Data-parallel training step of the PRODUCT (FusedTrainer: flat gradient slab with the [SSE, count] tail,
one all-reduce, rescale inside hcg_adam_step) against the single-rank step on the concatenated batch —
the reference's step is `utils/utils_model.py:55-70` on ONE device; sharding it must not change it.

* two emulated ranks on one GPU: two trainers, their slabs summed by hand (fp32 add = what a 2-rank NCCL
  sum all-reduce computes), both apply the identical update;
* two real ranks over NCCL when at least two GPUs are visible (skipped otherwise);
* hcg_adam_step alone against torch.optim.Adam on identical gradients.
"""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hcatgnet_b200 as p
    return p


def _fresh(pkg, sd=None, dev="cuda"):
    m = pkg.make_network("GCN", pkg.default_options(), 25)
    if sd is not None:
        m.load_state_dict(sd)
    return m.to(dev).train()


def _shards(pkg, n_graphs=600, cut=217, seed=5):
    from hcatgnet_b200.synth import synth_batch_numpy
    d = synth_batch_numpy(n_graphs, 25, seed)
    ptr = d["ptr"]
    graphs, e0 = [], 0
    gid = np.searchsorted(ptr, d["edge_index"][0], side="right") - 1
    counts = np.bincount(gid, minlength=n_graphs)
    for g in range(n_graphs):
        n0, n1, e1 = ptr[g], ptr[g + 1], e0 + counts[g]
        graphs.append({"x": torch.from_numpy(d["x"][n0:n1]), "edge_index": torch.from_numpy(d["edge_index"][:, e0:e1] - n0),
                       "y": torch.tensor([d["y"][g]]), "idx": g})
        e0 = e1
    mk = pkg.Batch.from_data_list
    return mk(graphs), mk(graphs[:cut]), mk(graphs[cut:])     # uneven shards on purpose


def test_two_emulated_ranks_equal_the_single_rank_step(pkg):
    """Sum of the two shards' [grad | SSE | count] slabs -> hcg_adam_step == one step on the whole batch.
    Replicas stay bit-identical; against the single-rank step the only difference is fp32 summation order
    (per-CTA partial sums over a different split of the graphs): a few ulp of the gradient."""
    whole, sh_a, sh_b = _shards(pkg)
    sd0 = {k: v.clone() for k, v in _fresh(pkg).state_dict().items()}
    one = pkg.FusedTrainer(_fresh(pkg, sd0))
    ra, rb = pkg.FusedTrainer(_fresh(pkg, sd0)), pkg.FusedTrainer(_fresh(pkg, sd0))
    for it in range(4):
        l1 = one.step(whole.to("cuda"))
        ra.accumulate(sh_a.to("cuda"))
        rb.accumulate(sh_b.to("cuda"))
        total = ra.grad + rb.grad                       # the 2-rank sum all-reduce
        assert float(total[-1]) == 600.0                # counts add up
        if it == 0:
            # same parameters on both sides: the gradient of the SUM of squared errors is additive over shards, the
            # only difference is the fp32 summation order (later steps start from parameters that already differ in
            # their last bits, which Adam and the network amplify: they are compared on the scale of the step below)
            assert torch.allclose(total[:-2], one.grad[:-2], rtol=2e-5, atol=2e-5 * float(one.grad[:-2].abs().max()))
            assert torch.allclose(total[-2], one.grad[-2], rtol=1e-5)
        ra.grad.copy_(total); rb.grad.copy_(total)
        la, lb = ra.apply(), rb.apply()
        assert torch.equal(ra.flat, rb.flat) and torch.equal(la, lb)        # replicas: identical bits
        assert abs(float(la) - float(l1)) <= 1e-3 * max(1.0, abs(float(l1)))
        if it == 0:
            # Adam with the reference's eps = 1e-9 turns a last-bit difference of a near-zero gradient into a visible
            # fraction of one lr-sized step (update = lr * m / (sqrt(v) + eps)), and the next steps amplify it
            # (measured: 1e-5 after step 1, 2e-4 after step 2, 5e-3 after step 4 - the same would separate two runs of
            # the reference on different machines).  Only the first update is a like-for-like comparison.
            assert float((ra.flat - one.flat).abs().max()) <= 1e-2 * 0.01
    assert float(ra.checksum()) == float(rb.checksum())


def _nccl_worker(rank, world, port, sd0, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist

    import hcatgnet_b200 as pkg
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    _, sh_a, sh_b = _shards(pkg)
    model = pkg.make_network("GCN", pkg.default_options(), 25)
    model.load_state_dict(sd0)
    tr = pkg.FusedTrainer(model.to(f"cuda:{rank}").train())
    losses, first = [], None
    for it in range(4):
        losses.append(float(tr.step((sh_a if rank == 0 else sh_b).to(f"cuda:{rank}"))))
        if it == 0:
            first = tr.flat.cpu().clone()
    ret[rank] = (losses, tr.flat.cpu(), float(tr.checksum()), first)
    dist.destroy_process_group()


def test_two_nccl_ranks_equal_the_single_rank_step(pkg):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    sd0 = {k: v.cpu().clone() for k, v in _fresh(pkg).state_dict().items()}
    ret = mp.Manager().dict()
    mp.spawn(_nccl_worker, args=(2, port, sd0, ret), nprocs=2, join=True)
    whole, _, _ = _shards(pkg)
    one = pkg.FusedTrainer(_fresh(pkg, sd0))
    ref, first = [], None
    for it in range(4):
        ref.append(float(one.step(whole.to("cuda"))))
        if it == 0:
            first = one.flat.cpu().clone()
    (l0, f0, c0, p0), (l1, f1, c1, _) = ret[0], ret[1]
    assert float((p0 - first).abs().max()) <= 1e-2 * 0.01                     # first update: like for like (see above)
    assert torch.equal(f0, f1) and c0 == c1 and l0 == l1                     # replicas: identical bits
    assert np.allclose(l0, ref, rtol=1e-3)


@pytest.mark.parametrize("with_tail", [False, True])
def test_adam_kernel_matches_torch_adam(pkg, with_tail):
    """hcg_adam_step vs torch.optim.Adam(lr 0.01, eps 1e-9: the reference's optimizer, model/networks.py:36-44)
    on identical gradients for 20 steps.  With the tail the kernel first scales by 1 / (2 count rmse)."""
    from hcatgnet_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(11)
    n = 14145
    p0 = torch.randn(n, generator=g)
    p_ref = p0.clone().cuda().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=0.01, eps=1e-9)
    p, m, v = p0.clone().cuda(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    loss_out = torch.zeros(1, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for step in range(1, 21):
        grad = (torch.randn(n, generator=g) * (10.0 ** float(torch.randint(-3, 2, (1,), generator=g)))).cuda()
        tail, ref_grad = None, grad
        if with_tail:
            sse, cnt = float(torch.rand(1, generator=g)) * 500 + 1.0, float(torch.randint(8, 5000, (1,), generator=g))
            tail = torch.tensor([sse, cnt], device="cuda")
            rmse = torch.sqrt(tail[0] / tail[1])
            ref_grad = grad * (1.0 / (2.0 * tail[1] * rmse))
        p_ref.grad = ref_grad.clone()
        opt.step()
        _lib.check(lib.hcg_adam_step(p.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), n,
                                     tail.data_ptr() if with_tail else None, 0.01, 0.9, 0.999, 1e-9, step, loss_out.data_ptr(), st))
        if with_tail:
            assert abs(float(loss_out) - float(rmse)) <= 1e-6 * float(rmse)
        err = float((p - p_ref.detach()).abs().max())
        assert err <= 1e-6 * max(1.0, float(p_ref.detach().abs().max())), (step, err)
    st_ref = opt.state[p_ref]
    # (the moments hold differences of large terms: absolute tolerance relative to their largest entry)
    assert float((m - st_ref["exp_avg"]).abs().max()) <= 2e-6 * float(st_ref["exp_avg"].abs().max())
    assert float((v - st_ref["exp_avg_sq"]).abs().max()) <= 2e-6 * float(st_ref["exp_avg_sq"].abs().max())


def test_trainer_survives_model_to_and_checkpoints_its_moments(pkg):
    """predict_network of the reference does model.to('cpu') (utils/utils_model.py:83): the parameters leave the
    slab.  The next step must notice (re-bind after model.to('cuda')), and the Adam moments must round-trip."""
    _, sh_a, _ = _shards(pkg)
    model = _fresh(pkg)
    tr = pkg.FusedTrainer(model)
    tr.step(sh_a.to("cuda"))
    model.to("cpu")
    with pytest.raises(RuntimeError):
        tr.step(sh_a.to("cuda"))
    model.to("cuda")
    before = [p.detach().clone() for p in model.parameters()]
    tr.step(sh_a.to("cuda"))
    assert all(p.data_ptr() == tr.flat.data_ptr() + 4 * o for p, o in zip(tr.params, tr._offsets))
    assert any(not torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))   # forward sees the update
    sd = tr.state_dict()
    tr2 = pkg.FusedTrainer(_fresh(pkg, model.state_dict()))
    tr2.load_state_dict(sd)
    la, lb = tr.step(sh_a.to("cuda")), tr2.step(sh_a.to("cuda"))
    assert torch.equal(la, lb) and torch.equal(tr.flat, tr2.flat)
