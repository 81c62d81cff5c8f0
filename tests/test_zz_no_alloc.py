"""gw_verify_pairs_noalloc (include/gw_align.h): the pair verifier for callers that cannot allocate device memory — DP history in a static
device array, device pointers in and out, nothing allocated by the call. No counterpart in the reference; checked against the existing
entry point gw_verify_pairs (itself checked against the oracle in test_gpu_parity.py), bit for bit, on two calls in a row that share the
scratch array. Collected last on purpose (the file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STRIDE = 40
MAX_BYTES = 64 << 20   # GW_VERIFY_PAIRS_NOALLOC_MAX_BYTES


def _pairs(n, seed, lengths):
    """(window, query) strings shaped like the aligner's verification requests: query = a genome piece with 0..6 edits, window = piece +- k"""
    from genome_weaver_align_b200 import synth
    rng = np.random.default_rng(seed)
    g = synth.random_genome(200_000, seed, n_frac=0.002)
    out = []
    for i in range(n):
        m = int(lengths[i % len(lengths)])
        s = int(rng.integers(10, len(g) - m - 10))
        q = bytearray(g[s:s + m].tobytes())
        for _ in range(int(rng.integers(0, 7))):
            p, kind = int(rng.integers(0, len(q))), int(rng.integers(0, 3))
            if kind == 0:
                q[p] = b"ACGT"[int(rng.integers(0, 4))]
            elif kind == 1:
                q.insert(p, b"ACGT"[int(rng.integers(0, 4))])
            elif len(q) > 1:
                del q[p]
        q = bytes(q[:256])
        if i % 97 == 0:   # an unrelated query: no alignment within the band
            q = synth.random_genome(len(q), seed + i).tobytes()
        out.append((g[s - 6:s + m + 6].tobytes().decode(), q.decode()))
    return out


def _existing(L, pairs, k):
    refs, ro = L.concat([r for r, _ in pairs])
    qs, qo = L.concat([q for _, q in pairs])
    n = len(pairs)
    res = np.zeros((n, 4), dtype=np.int32)
    cig = np.zeros((n, STRIDE), dtype=np.uint32)
    L.check(L.lib().gw_verify_pairs(0, L.ptr(refs), L.ptr(ro), L.ptr(qs), L.ptr(qo), n, k, L.ptr(res), L.ptr(cig), STRIDE))
    return res, cig


def _noalloc(L, torch, pairs, k, bufs):
    """through the C-ABI with device buffers made by the CALLER (torch here, allocated before the call)"""
    refs, ro = L.concat([r for r, _ in pairs])
    qs, qo = L.concat([q for _, q in pairs])
    n = len(pairs)
    d_refs, d_ro, d_qs, d_qo, d_res, d_cig = bufs
    d_refs[:len(refs)].copy_(torch.from_numpy(np.frombuffer(refs, dtype=np.uint8).copy()))
    d_qs[:len(qs)].copy_(torch.from_numpy(np.frombuffer(qs, dtype=np.uint8).copy()))
    d_ro[:n + 1].copy_(torch.from_numpy(np.asarray(ro, dtype=np.uint64).view(np.int64).copy()))
    d_qo[:n + 1].copy_(torch.from_numpy(np.asarray(qo, dtype=np.uint64).view(np.int64).copy()))
    d_res.fill_(-7)       # outputs start as garbage: the entry point must write every word it owns
    d_cig.fill_(-1)
    torch.cuda.synchronize()
    max_ref, max_q = max(len(r) for r, _ in pairs), max(len(q) for _, q in pairs)
    free_before = torch.cuda.mem_get_info()[0]
    rc = L.lib().gw_verify_pairs_noalloc(0, None, d_refs.data_ptr(), d_ro.data_ptr(), d_qs.data_ptr(), d_qo.data_ptr(), n, max_ref, max_q, k,
                                         d_res.data_ptr(), d_cig.data_ptr(), STRIDE)
    L.check(rc)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    res = d_res[:n * 4].cpu().numpy().reshape(n, 4)
    cig = d_cig[:n * STRIDE].cpu().numpy().view(np.uint32).reshape(n, STRIDE)
    return res, cig, free_before - free_after


def test_no_alloc_verifier_matches_the_existing_entry_point_twice():
    import torch
    from genome_weaver_align_b200 import _lib as L
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    cap_n, cap_bytes = 6000, 6000 * 300
    bufs = (torch.empty(cap_bytes, dtype=torch.uint8, device=dev), torch.empty(cap_n + 1, dtype=torch.int64, device=dev),
            torch.empty(cap_bytes, dtype=torch.uint8, device=dev), torch.empty(cap_n + 1, dtype=torch.int64, device=dev),
            torch.empty(cap_n * 4, dtype=torch.int32, device=dev), torch.empty(cap_n * STRIDE, dtype=torch.int32, device=dev))
    # call 1: 3000 pairs of mixed lengths (1 to 4 DP blocks); call 2: other, fewer and shorter pairs over the scratch call 1 left behind;
    # call 3: call 1's pairs again
    first = _pairs(3000, 101, [36, 64, 65, 100, 128, 150, 200, 250])
    second = _pairs(2000, 202, [50, 100, 75])
    for call, (pairs, k) in enumerate(((first, 6), (second, 3), (first, 6))):
        want_res, want_cig = _existing(L, pairs, k)
        got_res, got_cig, grew = _noalloc(L, torch, pairs, k, bufs)
        # the first use in a process may load the CUDA module that holds the static array (that is when the GPU commits its 64 MiB);
        # from then on the memory in use on the device must not move across a call
        assert call == 0 or grew <= 0, "device memory in use grew by %d bytes across call %d" % (grew, call)
        assert np.array_equal(got_res, want_res)
        assert np.array_equal(got_cig, want_cig)
        assert int((want_res[:, 0] != 0).sum()) > len(pairs) // 2   # the comparison is over real alignments, not over empty results


def test_no_alloc_verifier_rejects_what_the_scratch_cannot_hold():
    import torch
    from genome_weaver_align_b200 import _lib as L
    torch.cuda.set_device(0)
    z = torch.zeros(16, dtype=torch.int64, device="cuda:0")
    n = MAX_BYTES // ((106 + 2) * 2 * 16) + 1    # one pair more than the static array holds at 106 reference columns and 100-base queries
    rc = L.lib().gw_verify_pairs_noalloc(0, None, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), n, 106, 100, 3, z.data_ptr(), z.data_ptr(), STRIDE)
    assert rc == -1   # GW_E_ARG, nothing launched
    with pytest.raises(L.GwError, match="static scratch"):
        L.check(rc)
    # a pair longer than the caller stated is marked, not aligned, and does not touch memory beyond its share
    pairs = [("ACGTACGTACGTACGTACGT", "ACGTACGTAC"), ("ACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT", "ACGTACGTAC")]
    refs, ro = L.concat([r for r, _ in pairs])
    qs, qo = L.concat([q for _, q in pairs])
    d = [torch.from_numpy(np.frombuffer(b, dtype=np.uint8).copy()).cuda() for b in (refs, qs)]
    o = [torch.from_numpy(np.asarray(a, dtype=np.uint64).view(np.int64).copy()).cuda() for a in (ro, qo)]
    res = torch.full((8,), -7, dtype=torch.int32, device="cuda:0")
    cig = torch.full((2 * STRIDE,), -1, dtype=torch.int32, device="cuda:0")
    L.check(L.lib().gw_verify_pairs_noalloc(0, None, d[0].data_ptr(), o[0].data_ptr(), d[1].data_ptr(), o[1].data_ptr(), 2, 20, 10, 3, res.data_ptr(),
                                            cig.data_ptr(), STRIDE))
    torch.cuda.synchronize()
    r = res.cpu().numpy().reshape(2, 4)
    want, _ = _existing(L, pairs[:1], 3)
    assert list(r[0]) == list(want[0]) and r[1, 0] == -2 and not cig.cpu().numpy()[STRIDE:].any()
