"""The allocation-free cross-entropy forward (``bdl_xent_fwd_noalloc``), called through the C ABI.

Why it exists: some callers of the kernel library cannot get device memory when they call in — an inference server
that gave all of its HBM to one arena at start-up, a control loop that must not hit the device-wide stall of a
``cudaMalloc``, a plug-in living inside a runtime whose allocator a C caller cannot reach. The regular entry point
``bdl_xent_fwd`` makes its caller bring 3*B floats of scratch; the variant keeps that scratch in a fixed array inside the
library (at most 65536 rows) so such callers only pass inputs and outputs.

The file name sorts last on purpose: if this add-on ever breaks, a ``pytest -x`` run has already exercised everything
else. Each check makes two calls in a row, because the built-in scratch is shared between calls and keeps whatever the
previous call left in it.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_pair(lib, stream, logits, target, class_w, size_average):
    """Run the regular entry point (scratch from the caller) and the allocation-free one on the same inputs."""
    B, C = logits.shape
    is_bf16 = int(logits.dtype == torch.bfloat16)
    is_i64 = int(target.dtype == torch.int64)
    w_ptr = None if class_w is None else class_w.data_ptr()
    scratch = torch.empty(3 * B, dtype=torch.float32, device=logits.device)
    out_ref = torch.full((2,), -1.0, device=logits.device)
    out_new = torch.full((2,), -2.0, device=logits.device)
    rc = lib.bdl_xent_fwd(logits.data_ptr(), target.data_ptr(), w_ptr, scratch.data_ptr(), out_ref.data_ptr(), B, C,
                          logits.stride(0), is_bf16, is_i64, -1, int(size_average), stream)
    assert rc == 0, lib.bdl_last_error().decode()
    rc = lib.bdl_xent_fwd_noalloc(logits.data_ptr(), target.data_ptr(), w_ptr, out_new.data_ptr(), B, C,
                                  logits.stride(0), is_bf16, is_i64, -1, int(size_average), stream)
    assert rc == 0, lib.bdl_last_error().decode()
    torch.cuda.synchronize()
    return out_ref.cpu(), out_new.cpu()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    from bigdl_b200.ops import _lib

    lib = _lib.cuda_lib()
    assert lib is not None and hasattr(lib, "bdl_xent_fwd_noalloc")
    stream = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(11)
    dev = "cuda"

    # call 1: large fp32 batch, float (1-based) targets, unweighted mean
    lg1 = torch.randn(4096, 1000, device=dev) * 3
    tg1 = torch.randint(1, 1001, (4096,), device=dev).float()
    ref1, new1 = _call_pair(lib, stream, lg1, tg1, None, True)
    assert torch.allclose(new1, ref1, rtol=1e-6, atol=1e-6), (new1, ref1)
    plain = torch.nn.functional.cross_entropy(lg1, tg1.long() - 1)
    assert abs(float(new1[0]) - float(plain)) < 2e-3 * max(1.0, abs(float(plain)))

    # call 2, straight after: smaller bf16 batch with class weights, int64 targets, summed loss — the scratch still
    # holds call 1's rows beyond row 100, which must not leak into the result
    lg2 = (torch.randn(100, 37, device=dev) * 2).to(torch.bfloat16)
    tg2 = torch.randint(1, 38, (100,), device=dev)
    w2 = (torch.rand(37, device=dev) + 0.5).contiguous()
    ref2, new2 = _call_pair(lib, stream, lg2, tg2, w2, False)
    assert torch.allclose(new2, ref2, rtol=1e-6, atol=1e-6), (new2, ref2)
    plain2 = torch.nn.functional.cross_entropy(lg2.float(), tg2 - 1, weight=w2, reduction="sum")
    assert abs(float(new2[0]) - float(plain2)) < 2e-3 * max(1.0, abs(float(plain2)))


def test_noalloc_rejects_oversized_batch():
    from bigdl_b200.ops import _lib

    lib = _lib.cuda_lib()
    stream = torch.cuda.current_stream().cuda_stream
    lg = torch.zeros(8, 4, device="cuda")
    tg = torch.ones(8, device="cuda")
    out = torch.zeros(2, device="cuda")
    rc = lib.bdl_xent_fwd_noalloc(lg.data_ptr(), tg.data_ptr(), None, out.data_ptr(), 65537, 4, 4, 0, 0, -1, 1, stream)
    assert rc != 0
    assert "65536" in lib.bdl_last_error().decode()
    torch.cuda.synchronize()
