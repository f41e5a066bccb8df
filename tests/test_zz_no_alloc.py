"""hcg_gcn_forward_coo_static - the edge-list forward with its scratch in a file-scope device array (no workspace
argument, no allocation; added for callers that cannot call cudaMalloc, no counterpart in the reference) - against
hcg_gcn_forward_coo on the same inputs, through the C-ABI, on two calls in a row that share the array.
Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hcatgnet_b200 as p
    return p


def _call_pair(pkg, model, n_graphs, seed):
    """(out, emb) of the existing entry point and of the static one on the same synthetic batch."""
    from hcatgnet_b200 import _lib, ops
    from hcatgnet_b200.synth import synth_batch
    lib = _lib.load()
    b = synth_batch(n_graphs, 25, seed=seed).to("cuda")
    dims = model._dims()
    params = [p.detach() for p in model._flat_params()]
    ps = ops._params_struct(dims, params)
    N, E, B = int(b.x.shape[0]), int(b.edge_index.shape[1]), int(b.num_graphs)
    st = torch.cuda.current_stream().cuda_stream
    f32 = dict(dtype=torch.float32, device="cuda")
    out_w, emb_w = torch.empty(B, dims.C, **f32), torch.empty(B, 2 * dims.H, **f32)
    ws = torch.empty(lib.hcg_forward_coo_workspace_bytes(B, C.byref(ps)), dtype=torch.uint8, device="cuda")
    _lib.check(lib.hcg_gcn_forward_coo(b.x.data_ptr(), b.edge_index.data_ptr(), b.ptr.data_ptr(), b.edge_ptr.data_ptr(), N, E, B,
                                       1.0, C.byref(ps), out_w.data_ptr(), emb_w.data_ptr(), ws.data_ptr(), ws.numel(), st))
    flags_w = C.c_int32(-1)
    _lib.check(lib.hcg_forward_coo_status(ws.data_ptr(), B, st, C.byref(flags_w)))
    out_s, emb_s = torch.full((B, dims.C), float("nan"), **f32), torch.full((B, 2 * dims.H), float("nan"), **f32)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    flags_s = C.c_int32(-1)
    _lib.check(lib.hcg_gcn_forward_coo_static(b.x.data_ptr(), b.edge_index.data_ptr(), b.ptr.data_ptr(), b.edge_ptr.data_ptr(), N, E,
                                              B, 1.0, C.byref(ps), out_s.data_ptr(), emb_s.data_ptr(), C.byref(flags_s), st))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert flags_w.value == 0 and flags_s.value == 0
    return (out_w, emb_w), (out_s, emb_s), free0 - free1


def test_static_scratch_forward_matches_the_workspace_forward_twice_in_a_row(pkg):
    torch.manual_seed(2)
    model = pkg.make_network("GCN", pkg.default_options(), 25).cuda().eval()
    # a large batch first, then a small one: the second call finds the first call's work list in the shared array
    for call, (n_graphs, seed) in enumerate([(3000, 41), (77, 42)]):
        (out_w, emb_w), (out_s, emb_s), grew = _call_pair(pkg, model, n_graphs, seed)
        # same kernels on the same inputs: the tolerance of tests/test_gpu_edge_list.py for this path is equality
        assert torch.equal(emb_s, emb_w), call
        assert torch.equal(out_s, out_w), call
        assert torch.isfinite(out_s).all()
        if call > 0:        # (the first call of a process may load the module / set the kernel's shared-memory limit)
            assert grew <= 0, f"device memory in use grew by {grew} bytes across a call that must not allocate"


def test_static_scratch_forward_rejects_what_does_not_fit(pkg):
    from hcatgnet_b200 import _lib, ops
    lib = _lib.load()
    model = pkg.make_network("GCN", pkg.default_options(), 25).cuda().eval()
    ps = ops._params_struct(model._dims(), [p.detach() for p in model._flat_params()])
    dummy = torch.zeros(16, device="cuda")
    idx = torch.zeros(16, dtype=torch.int64, device="cuda")
    rc = lib.hcg_gcn_forward_coo_static(dummy.data_ptr(), idx.data_ptr(), idx.data_ptr(), idx.data_ptr(), 0, 0, 16385, 1.0,
                                        C.byref(ps), dummy.data_ptr(), dummy.data_ptr(), None,
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 3          # HCG_ERR_WORKSPACE (include/hcg.h)
    assert b"16384" in lib.hcg_last_error()
