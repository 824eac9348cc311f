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


# ------------------------------------------------------------------------------------------------ bucketed kernels
CM_COMPRESS, CM_PUBLISH_LOCAL, CM_CLIP_DEV = 512, 1024, 4096


def _bucket_launch(lib, grp_g16, grp_w16, world, r, b_off, b_len, s, l, grads32, master, gpart, mom, kind, hp, mode,
                   clip_scale=None, sumsq=None):
    from bigdl_b200.ops import _lib

    ls = torch.zeros(4, dtype=torch.int32, device="cuda")
    buf = ctypes.create_string_buffer(hp, len(hp))
    _lib.check(lib.bdl_bucket_rs_update(grp_g16, grp_w16, None, world, r, b_off, b_len, s, l,
                                        None if grads32 is None else grads32.data_ptr(),
                                        None if master is None else master.data_ptr(),
                                        None if gpart is None else gpart.data_ptr(),
                                        None if mom is None else mom.data_ptr(), None, kind,
                                        ctypes.cast(buf, ctypes.c_void_p), None, mode, 1, 1, None, 0, ls.data_ptr(),
                                        None if sumsq is None else sumsq.data_ptr(),
                                        None if clip_scale is None else clip_scale.data_ptr(), 0, 0, 16,
                                        torch.cuda.current_stream().cuda_stream), "bdl_bucket_rs_update")


@pytest.mark.parametrize("world", [2, 4, 8])
def test_bucketed_reduce_update_and_pull_virtual_ranks(world):
    """Two buckets of one flat vector, N virtual ranks on one GPU (phases launched rank by rank instead of the flag
    barriers): in-kernel compress, sub-slice reduce + SGD-momentum update, local publish, pulled all-gather; then the
    two-phase global-norm clip with the scale read from device memory."""
    from bigdl_b200.ops import _lib
    from bigdl_b200.optim.optim_method import pack_hp
    from bigdl_b200.parallel.bucketed import build_buckets
    from bigdl_b200.parallel.symmetric import LocalPeerGroup

    lib = _lib.cuda_lib()
    dev = torch.device("cuda")
    torch.manual_seed(1)
    P = 8 * 3000
    units = [(None, 0, 8 * 1100), (None, 8 * 1100, P - 8 * 1100)]
    buckets, ptot = build_buckets(units, world, cap=1, first_cap=1)
    assert len(buckets) == 2
    off_w = (P * 2 + 255) // 256 * 256
    grp = LocalPeerGroup(world, off_w * 2 + 256, dev)
    g16, w16 = grp.peer_ptr_array(0), grp.peer_ptr_array(off_w)
    grads = [torch.randn(P, device=dev) for _ in range(world)]
    master_full = torch.randn(P, device=dev)
    work = [g.clone() for g in grads]
    hp_kw = dict(lr=0.05, wd=1e-3, momentum=0.9, dampening=0.0, first_step=1)
    hp = pack_hp(gscale=1.0 / world, **hp_kw)
    parts = [torch.cat([master_full[b.slices[r][0]:b.slices[r][0] + b.slices[r][1]] for b in buckets]).clone()
             for r in range(world)]
    gparts = [torch.zeros(ptot[r], device=dev) for r in range(world)]
    moms = [torch.zeros(ptot[r], device=dev) for r in range(world)]
    for b in buckets:      # phase 0 on every rank, then phase 1 on every rank (what the begin barrier orders on real peers)
        for r in range(world):
            _bucket_launch(lib, g16, w16, world, r, b.off, b.length, b.slices[r][0], 0, work[r], None, None, None, 0, hp,
                           CM_COMPRESS)
        for r in range(world):
            s, l = b.slices[r]
            po = b.poff[r]
            _bucket_launch(lib, g16, w16, world, r, b.off, b.length, s, l, None, parts[r][po:po + l], gparts[r][po:po + l],
                           moms[r][po:po + l], 0, hp, CM_REDUCE | CM_WRITE_GRAD | CM_UPDATE | CM_PUBLISH_LOCAL)
    torch.cuda.synchronize()
    for r in range(world):
        assert float(work[r].abs().max()) == 0.0                  # zeroGradParameters fused into the compress phase
    ref_g = sum(g.to(torch.bfloat16).float() for g in grads) / world
    gd = ref_g + 1e-3 * master_full
    ref_w = master_full - 0.05 * gd
    for b in buckets:
        for r in range(world):
            s, l = b.slices[r]
            po = b.poff[r]
            assert torch.allclose(gparts[r][po:po + l], ref_g[s:s + l], rtol=1e-5, atol=1e-6)
            assert torch.allclose(parts[r][po:po + l], ref_w[s:s + l], rtol=1e-5, atol=1e-6)
            assert torch.allclose(moms[r][po:po + l], gd[s:s + l], rtol=1e-5, atol=1e-6)
            own = grp.view(r, torch.bfloat16, off_w, P)[s:s + l]
            assert torch.equal(own, parts[r][po:po + l].to(torch.bfloat16))          # only the owner's copy so far
    # pull: every rank fetches the others' sub-slices
    for r in range(world):
        for b in buckets:
            segs = [(q, b.slices[q][0], b.slices[q][1]) for q in range(world) if q != r and b.slices[q][1]]
            arr = (ctypes.c_longlong * (3 * len(segs)))(*[v for sgm in segs for v in sgm])
            flags = torch.zeros(64, dtype=torch.int32, device=dev)
            _lib.check(lib.bdl_pull_gather(w16, flags.data_ptr(), world, r, arr, len(segs), 2, b.index + 1, 1, None, 0, 8,
                                           torch.cuda.current_stream().cuda_stream), "bdl_pull_gather")
    torch.cuda.synchronize()
    full = ref_w.clone()
    for b in buckets:
        for r in range(world):
            s, l = b.slices[r]
            full[s:s + l] = parts[r][b.poff[r]:b.poff[r] + l]
    for r in range(world):
        assert torch.equal(grp.view(r, torch.bfloat16, off_w, P), full.to(torch.bfloat16))
    # phase 2 of global-norm clipping: update from the stored averaged gradient, scaled by a device-resident factor
    scale = torch.tensor([0.25], device=dev)
    hp2 = pack_hp(gscale=1.0 / world, lr=0.1)          # gscale must NOT be applied a second time
    r, b = 0, buckets[1]
    s, l = b.slices[r]
    po = b.poff[r]
    before = parts[r][po:po + l].clone()
    _bucket_launch(lib, g16, w16, world, r, b.off, b.length, s, l, None, parts[r][po:po + l], gparts[r][po:po + l], None, 0,
                   hp2, CM_UPDATE | CM_PUBLISH_LOCAL | CM_CLIP_DEV, clip_scale=scale)
    torch.cuda.synchronize()
    assert torch.allclose(parts[r][po:po + l], before - 0.1 * 0.25 * gparts[r][po:po + l], rtol=1e-5, atol=1e-6)


def test_peer_clip_scale_single_rank():
    from bigdl_b200.ops import _lib

    lib = _lib.cuda_lib()
    dev = torch.device("cuda")
    slots = torch.zeros(4, device=dev)
    flags = torch.zeros(8, dtype=torch.int32, device=dev)
    sumsq = torch.tensor([16.0], device=dev)
    scale, norm = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    sp = (ctypes.c_void_p * 1)(slots.data_ptr())
    fp = (ctypes.c_void_p * 1)(flags.data_ptr())
    epoch = torch.zeros(1, dtype=torch.int64, device=dev)
    for step, (ss, want) in enumerate([(16.0, 0.5), (1.0, 1.0)]):
        sumsq.fill_(ss)
        epoch.fill_(step)
        _lib.check(lib.bdl_peer_clip_scale(sp, fp, 1, 0, 1, 1, epoch.data_ptr(), 0, sumsq.data_ptr(), 2.0,
                                           scale.data_ptr(), norm.data_ptr(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        assert abs(float(scale) - want) < 1e-6 and abs(float(norm) - ss ** 0.5) < 1e-6 and float(sumsq) == 0.0
