"""The tensor-core building block (split-TF32 tcgen05.mma, A operand in tensor memory) against an
fp64 matmul: the accuracy the GCN parity bound (1e-5 relative) relies on."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("K", [64, 32, 25, 8, 57])
def test_tc_split_tf32_tile_product(K):
    from hcatgnet_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(K)
    A = (torch.randn(128, K, generator=g) * 3).cuda()
    W = torch.randn(64, K, generator=g).cuda()
    D = torch.full((128, 64), float("nan"), device="cuda")
    _lib.check(lib.hcg_selftest_tc_gemm(A.data_ptr(), W.data_ptr(), K, D.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    ref = A.double().cpu() @ W.double().cpu().t()
    scale = (A.double().abs().cpu() @ W.double().abs().cpu().t()).clamp(min=1e-30)   # condition-free error measure
    err = ((D.double().cpu() - ref).abs() / scale).max().item()
    print(f"K={K}: max |D - ref| / (|A| |W|^T) = {err:.3e}")
    # fp32-class accuracy (a plain TF32 pass would sit at 2^-11 ~ 5e-4)
    assert err <= 5e-7, err


@pytest.mark.parametrize("K", [128, 80, 41, 8])
def test_tc_split_tf32_xty_product(K):
    """G = X^T Y with the node index as the reduction index (MN-major shared-memory operands): the building block
    of the weight gradients g_W = gT^T h on the tensor cores."""
    from hcatgnet_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(100 + K)
    X = (torch.randn(128, 64, generator=g) * 2).cuda()
    Y = torch.randn(128, 64, generator=g).cuda()
    G = torch.full((64, 64), float("nan"), device="cuda")
    _lib.check(lib.hcg_selftest_tc_xty(X.data_ptr(), Y.data_ptr(), K, G.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    ref = X[:K].double().cpu().t() @ Y[:K].double().cpu()
    scale = (X[:K].double().abs().cpu().t() @ Y[:K].double().abs().cpu()).clamp(min=1e-30)
    err = ((G.double().cpu() - ref).abs() / scale).max().item()
    print(f"K={K}: max |G - ref| / (|X|^T |Y|) = {err:.3e}")
    assert err <= 1e-6, err
