"""GPU: pb_linear_fwd_noalloc -- the dense-layer forward for callers that cannot allocate device memory.

Some hosts of this library may not create device allocations where they call it (their own arena owns the whole
device heap, or a cudaMalloc's device-wide synchronisation is not acceptable in their serving loop, or they are
plugins allowed to launch kernels only).  For them the split-K scratch of the dense layer lives in a fixed 4 MiB
array inside the library.  This test drives the entry point through the C ABI exactly as such a caller would
(plain pointers, no workspace argument) and holds it to the ordinary entry point, pb_linear_fwd, on two calls in
a row that share the built-in scratch; it also checks that a problem beyond the documented limit is refused.
Collected last on purpose (file name), so a failure here cannot hide the rest of a `-x` run.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_and_noalloc(lib, M, N, K, act, seed):
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(M, K, generator=g).cuda()
    W = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    b = torch.randn(N, generator=g).cuda()
    st = torch.cuda.current_stream().cuda_stream
    # ordinary entry point, with caller-provided scratch
    ws = torch.empty(lib.pb_workspace_bytes(b"linear_fwd", M, N, K), dtype=torch.uint8, device="cuda")
    Z0, Y0 = torch.empty(M, N, device="cuda"), torch.empty(M, N, device="cuda")
    assert lib.pb_linear_fwd(X.data_ptr(), W.data_ptr(), b.data_ptr(), Z0.data_ptr(), Y0.data_ptr(), M, N, K, act,
                             ws.data_ptr(), ws.numel(), st) == 0, lib.pb_last_error()
    # the variant without any scratch argument
    Z1, Y1 = torch.full((M, N), float("nan"), device="cuda"), torch.full((M, N), float("nan"), device="cuda")
    rc = lib.pb_linear_fwd_noalloc(X.data_ptr(), W.data_ptr(), b.data_ptr(), Z1.data_ptr(), Y1.data_ptr(), M, N, K, act, st)
    assert rc == 0, lib.pb_last_error()
    torch.cuda.synchronize()
    ref = torch.nn.functional.gelu(X.double() @ W.double().t() + b.double()) if act == 1 else X.double() @ W.double().t() + b.double()
    return (Z0, Y0), (Z1, Y1), ref


def test_noalloc_matches_pb_linear_fwd_on_two_calls_in_a_row():
    import proteinbert_b200 as pb
    lib = pb._lib.lib()
    # call 1: the annotation -> global layer of the default model (split-K: uses the built-in scratch fully)
    # call 2: a smaller problem right after it, which finds the first call's partial sums still in the array
    for (M, N, K, act, seed) in [(64, 512, 8943, 1, 0), (16, 128, 3001, 1, 1)]:
        (Z0, Y0), (Z1, Y1), ref = _ref_and_noalloc(lib, M, N, K, act, seed)
        # same kernels, same split: the two entry points must agree to fp32 round-off (tests' 1e-4 bound is loose here)
        assert torch.allclose(Z1, Z0, rtol=1e-5, atol=1e-6) and torch.allclose(Y1, Y0, rtol=1e-5, atol=1e-6)
        err = float((Y1.double() - ref).norm() / ref.norm())
        assert err <= 1e-4, err


def test_noalloc_refuses_sizes_beyond_its_limit():
    import proteinbert_b200 as pb
    lib = pb._lib.lib()
    M, N, K = 64, 1024, 64            # 65536 output elements > 32768
    X, W = torch.zeros(M, K, device="cuda"), torch.zeros(N, K, device="cuda")
    Y = torch.zeros(M, N, device="cuda")
    rc = lib.pb_linear_fwd_noalloc(X.data_ptr(), W.data_ptr(), None, None, Y.data_ptr(), M, N, K, 0,
                                   torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b"exceeds" in lib.pb_last_error()
