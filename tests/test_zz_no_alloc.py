"""GPU: vsk_pair_dp_static, the pair-DP entry point that allocates nothing.

It serves callers that cannot allocate device memory at the moment they need an
alignment (the GPU belongs to their own pool allocator, or they sit in a section
where cudaMalloc's device-wide synchronisation is unacceptable, or their memory
budget is fixed): inputs and outputs are device buffers the caller already owns
(here: torch tensors), all scratch lives in a fixed array inside libvsk.so.

Checked here, through the C ABI: its results equal those of the existing
vsk_pair_dp entry point bit for bit (integer work, no tolerance) and those of the CPU
oracle, on two calls in a row (the shared scratch keeps the first call's contents),
and oversize requests or a band that vsk_band_params would not give are refused with
a message.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(seed, n, maxlen):
    from vntrseek_b200 import synth
    rng = np.random.default_rng(seed)
    profs, li, ti, mode = [], [], [], []
    for k in range(n):
        L = int(rng.integers(1, maxlen + 1))
        a = synth._column_ids(rng, rng.integers(0, 4, L))
        if k % 3 == 0:
            b = rng.integers(0, 1001, int(rng.integers(max(1, L // 2), 2 * L + 1)))
            a = rng.integers(0, 1001, L)
        elif L >= 7:
            b = synth._mutate_profile(rng, a, 0.05, 0.03)
        else:
            b = a.copy()
        profs += [a, b]
        nb_ok = max(len(a), len(b)) <= 2 * min(len(a), len(b))
        for m in ((0, 1) if nb_ok else (0,)):
            li.append(2 * k)
            ti.append(2 * k + 1)
            mode.append(m)
    return profs, np.array(li, np.int32), np.array(ti, np.int32), np.array(mode, np.uint8)


def _run_static(ctx, profs, li, ti, mode):
    import torch

    from vntrseek_b200 import pack_profiles
    from vntrseek_b200._lib import DpDesc, lib
    off, ids = pack_profiles(profs)
    d_ids = torch.from_numpy(ids.astype(np.int16)).cuda()       # caller-owned device input
    d_out = torch.full((len(li), 4), -7, dtype=torch.int32, device="cuda")  # caller-owned device output
    desc = (DpDesc * len(li))()
    for k in range(len(li)):
        l1 = int(off[li[k] + 1] - off[li[k]])
        l2 = int(off[ti[k] + 1] - off[ti[k]])
        sv, br = C.c_double(0.0), C.c_int32(-1)
        if mode[k]:
            assert lib().vsk_band_params(l1, l2, 88, C.byref(sv), C.byref(br)) == 0
        desc[k] = DpDesc(int(off[li[k]]), int(off[ti[k]]), l1, l2, br.value, k, sv.value)
    torch.cuda.synchronize()
    rc = lib().vsk_pair_dp_static(ctx.h, d_ids.data_ptr(), len(ids), d_ids.data_ptr(), len(ids), desc, len(li),
                                  d_out.data_ptr())
    return rc, d_out.cpu().numpy()


def test_static_matches_pair_dp_twice(ctx, oracle):
    for seed, n, maxlen in ((21, 200, 300), (22, 40, 1500)):   # two calls in a row, second with wide pairs
        profs, li, ti, mode = _batch(seed, n, maxlen)
        ps = ctx.profiles(profs)
        want = ctx.pair_dp(ps, ps, li, ti, mode, 88)
        ps.close()
        rc, got = _run_static(ctx, profs, li, ti, mode)
        assert rc == 0
        assert (got == want).all()
        for k in range(0, len(li), 7):    # and both against the CPU oracle, not only against each other
            assert tuple(int(x) for x in got[k]) == oracle.dp(profs[li[k]], profs[ti[k]], int(mode[k]), 88), k
    # and the first batch again: nothing left over from the wide one disturbs it
    profs, li, ti, mode = _batch(21, 200, 300)
    ps = ctx.profiles(profs)
    want = ctx.pair_dp(ps, ps, li, ti, mode, 88)
    ps.close()
    rc, got = _run_static(ctx, profs, li, ti, mode)
    assert rc == 0 and (got == want).all()


def test_static_refuses_oversize(ctx):
    from vntrseek_b200._lib import lib
    rng = np.random.default_rng(3)
    profs = [rng.integers(0, 1001, 2049), rng.integers(0, 1001, 50)]
    rc, got = _run_static(ctx, profs, np.array([0], np.int32), np.array([1], np.int32), np.array([0], np.uint8))
    assert rc != 0 and (got == -7).all()          # refused, output untouched
    assert b"2048" in lib().vsk_last_error(ctx.h)
    # a band that is not vsk_band_params' for these lengths is refused as well
    from vntrseek_b200 import pack_profiles
    import torch
    from vntrseek_b200._lib import DpDesc
    off, ids = pack_profiles([profs[1], profs[1]])
    d_ids = torch.from_numpy(ids.astype(np.int16)).cuda()
    d_out = torch.zeros((1, 4), dtype=torch.int32, device="cuda")
    bad = (DpDesc * 1)(DpDesc(0, 50, 50, 50, 9, 0, 0.5))
    torch.cuda.synchronize()
    assert lib().vsk_pair_dp_static(ctx.h, d_ids.data_ptr(), len(ids), d_ids.data_ptr(), len(ids), bad, 1,
                                    d_out.data_ptr()) != 0
    assert b"vsk_band_params" in lib().vsk_last_error(ctx.h)
    # rows are the limit, not columns: the transposed pair is fine
    rc, got = _run_static(ctx, profs, np.array([1], np.int32), np.array([0], np.int32), np.array([0], np.uint8))
    assert rc == 0 and (got != -7).all()
