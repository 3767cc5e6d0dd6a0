"""clb_match_2nn_noalloc: the allocation-free 2-NN entry point (static device scratch, no cudaMalloc
-- no counterpart in the reference) against the existing clb_match_2nn_dev on the same device buffers, two calls
in a row (the shared scratch array keeps the first call's contents), bit-exact as every integer result in this suite; plus the
oracle on the first problem and the size-limit error.  Named test_zz_* so that pytest collects it last."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from coloc_b200 import synth   # noqa: E402

pytestmark = pytest.mark.gpu


def _problem(nq, nt, seed):
    T = synth.random_descriptors(nt, seed); Q, _ = synth.planted_queries(T, nq, seed + 1)
    if nt > 8:
        T[nt - 1] = T[5]; T[nt // 2] = T[5]; Q[0] = T[5]          # duplicated rows: the first one must win
    return Q, T


def test_noalloc_equals_existing_entry_point_two_calls_in_a_row(clb, oracle):
    import torch
    clb.set_param("match_tc", 0)                                  # the existing entry on the same (POPC) kernels; results are path-independent anyway
    try:
        m = clb.Matcher(maxkp=4096)
        st = torch.cuda.current_stream().cuda_stream
        # second problem smaller than the first: its slots of the shared scratch still hold the first call's keys
        for k, (nq, nt, thr) in enumerate(((3000, 20011, 40), (700, 1500, 60), (1, 1, 0), (257, 5, 40))):
            Q, T = _problem(nq, nt, 11 + 7 * k)
            dQ = torch.from_numpy(Q.view(np.int64)).cuda(); dT = torch.from_numpy(T.view(np.int64)).cuda()
            want = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
            got = torch.full((3, nq), -7, dtype=torch.int32, device="cuda")
            m.match_dev(dQ, nq, dT, nt, thr, want[0], want[1], want[2], stream=st)
            clb.match_2nn_noalloc(dQ, nq, dT, nt, thr, got[0], got[1], got[2], stream=st)
            clb.match_2nn_noalloc(dQ, nq, dT, nt, thr, got[0], got[1], got[2], stream=st)     # and again, straight behind it
            torch.cuda.synchronize()
            assert torch.equal(got, want), (k, int((got != want).sum()))
            if k == 0:
                ref = oracle.match_2nn(Q, T, thr)
                assert all(np.array_equal(got[i].cpu().numpy(), ref[i]) for i in range(3))
        m.close()
    finally:
        clb.set_param("match_tc", -1)


def test_noalloc_against_tensor_core_entry_and_limits(clb):
    import torch
    # default dispatch of the existing entry (tensor-core kernel at this size) gives the same lists
    nq, nt = 5000, 5000
    Q, T = _problem(nq, nt, 91)
    dQ = torch.from_numpy(Q.view(np.int64)).cuda(); dT = torch.from_numpy(T.view(np.int64)).cuda()
    m = clb.Matcher(maxkp=8192)
    want = torch.zeros((3, nq), dtype=torch.int32, device="cuda"); got = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    m.match_dev(dQ, nq, dT, nt, 40, want[0], want[1], want[2], stream=st)
    clb.match_2nn_noalloc(dQ, nq, dT, nt, 40, got[0], None, None, stream=st)                 # optional outputs left out
    clb.match_2nn_noalloc(dQ, nq, dT, nt, 40, got[0], got[1], got[2], stream=st)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    # empty DB: sentinels of CUDAK2NN.cu:54, no match
    clb.match_2nn_noalloc(dQ, 10, dT, 0, 40, got[0], got[1], got[2], stream=st)
    torch.cuda.synchronize()
    assert got[0, :10].tolist() == [-1] * 10 and got[1, :10].tolist() == [100000] * 10 and got[2, :10].tolist() == [200000] * 10
    # sizes beyond the static scratch are refused with a message that names the limit; nothing is launched
    with pytest.raises(clb.ClbError) as e:
        clb.match_2nn_noalloc(dQ, 32769, dT, nt, 40, got[0])
    assert e.value.code == 1 and "32768" in str(e.value)
    with pytest.raises(clb.ClbError):
        clb.match_2nn_noalloc(dQ.data_ptr() + 8, 10, dT, nt, 40, got[0])                      # misaligned rows
    m.close()
