"""Fused NVLink reduce-scatter + update + all-gather kernel, exercised on ONE GPU with N virtual ranks whose buffers
all live on the device (barriers off; launch order gives the ordering the peer flags give across GPUs)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

CM_REDUCE, CM_WRITE_GRAD, CM_UPDATE, CM_PUBLISH, CM_PARITY, CM_SUMSQ, CM_DIVIDE = 1, 2, 4, 8, 64, 128, 256


def _trunc(t):
    return ((t.contiguous().view(torch.int32) >> 16) << 16).view(torch.float32)


def _run(world, P, parity, kind_hp):
    from bigdl_b200.ops import _lib
    from bigdl_b200.optim.optim_method import pack_hp
    from bigdl_b200.parallel.all_reduce_parameter import aligned_partition
    from bigdl_b200.parallel.symmetric import LocalPeerGroup

    lib = _lib.cuda_lib()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    off_w = (P * 2 + 255) // 256 * 256
    grp = LocalPeerGroup(world, off_w * 2 + 256, dev)
    grads = [torch.randn(P, device=dev) for _ in range(world)]
    master_full = torch.randn(P, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for r in range(world):
        g = grads[r].clone()
        _lib.check(lib.bdl_compress_grads(g.data_ptr(), grp.view(r, torch.int16, 0, P).data_ptr(), P, int(parity), 1, st))
        assert float(g.abs().max()) == 0.0        # fused zeroGradParameters
    g16 = grp.peer_ptr_array(0)
    w16 = grp.peer_ptr_array(off_w)
    masters, gparts = [], []
    kind, hp_kwargs = kind_hp
    for r in range(world):
        s, l = aligned_partition(P, world, r)
        m = master_full[s:s + l].clone()
        gp = torch.zeros(l, device=dev)
        mom = torch.zeros(l, device=dev)
        hp = pack_hp(gscale=float(world) if parity else 1.0 / world, **hp_kwargs)
        buf = ctypes.create_string_buffer(hp, len(hp))
        mode = CM_REDUCE | CM_WRITE_GRAD | CM_UPDATE | CM_PUBLISH | (CM_PARITY | CM_DIVIDE if parity else 0)
        _lib.check(lib.bdl_fused_rs_update(g16, w16, None, world, r, s, l, m.data_ptr(), gp.data_ptr(), mom.data_ptr(),
                                           None, kind, ctypes.cast(buf, ctypes.c_void_p), None, mode, 1, None, None, 0,
                                           None, st), "bdl_fused_rs_update")
        masters.append((s, l, m))
        gparts.append(gp)
    torch.cuda.synchronize()
    return grp, grads, master_full, masters, gparts, off_w


@pytest.mark.parametrize("world", [2, 4, 8])
def test_parity_mode_is_bit_exact(world):
    P = 8 * 1237
    grp, grads, master_full, masters, gparts, off_w = _run(world, P, True, (0, dict(lr=0.1)))
    acc = _trunc(grads[0])
    for r in range(1, world):
        acc = _trunc(acc + _trunc(grads[r]))
    ref_g = acc / world
    for (s, l, m), gp in zip(masters, gparts):
        assert torch.equal(gp, ref_g[s:s + l])
        ref_w = master_full[s:s + l] - 0.1 * ref_g[s:s + l]
        assert torch.allclose(m, ref_w, rtol=1e-6, atol=1e-7)
        pub = _trunc(m)
        for r in range(world):       # every virtual rank received the owner's truncated slice
            got = grp.view(r, torch.bfloat16, off_w, P)[s:s + l].float()
            assert torch.equal(got, pub)


@pytest.mark.parametrize("world", [2, 8])
def test_fast_mode_sgd_momentum(world):
    P = 8 * 4096
    hp = dict(lr=0.05, wd=1e-3, momentum=0.9, dampening=0.0, first_step=1)
    grp, grads, master_full, masters, gparts, off_w = _run(world, P, False, (0, hp))
    ref_g = sum(g.to(torch.bfloat16).float() for g in grads) / world
    for (s, l, m), gp in zip(masters, gparts):
        assert torch.allclose(gp, ref_g[s:s + l], rtol=1e-5, atol=1e-6)
        g = ref_g[s:s + l] + 1e-3 * master_full[s:s + l]
        ref_w = master_full[s:s + l] - 0.05 * g
        assert torch.allclose(m, ref_w, rtol=1e-5, atol=1e-6)
        for r in range(world):
            got = grp.view(r, torch.bfloat16, off_w, P)[s:s + l]
            assert torch.equal(got, m.to(torch.bfloat16))
