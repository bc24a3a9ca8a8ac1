"""The allocation-free entry point sc_exclusive_scan_i32_noalloc (include/scb200.h; no counterpart in the
reference) against the existing sc_exclusive_scan_i32, called
through the C-ABI.  Integer work: the bar is bit-exact, as for every index computation in this repository.

Collected last on purpose (file name).  Passes alone in a fresh process and
after the other tests: every buffer is allocated BEFORE the calls under test, and the test checks that the
calls themselves leave torch's allocator untouched."""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(n, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    # tile counts of the binning are small non-negative integers with many zeros; add large values so that
    # the int64 accumulation is exercised (the total passes 2^31)
    x = torch.randint(0, 7, (n,), generator=g, dtype=torch.int32)
    x[torch.randint(0, n, (max(n // 64, 1),), generator=g)] = 0
    x[torch.randint(0, n, (max(n // 1000, 1),), generator=g)] = 2 ** 30
    return x.to(dev)


def test_noalloc_scan_equals_workspace_scan_twice_in_a_row():
    from street_crafter_b200._lib import check, lib

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    st = torch.cuda.current_stream().cuda_stream
    max_n = int(lib.sc_exclusive_scan_i32_noalloc_max_n())
    assert max_n == 1 << 24

    # two calls in a row, the second SMALLER than the first so that it runs over the first call's left-overs
    # in the shared scratch; then the maximum size and two ragged sizes around a block boundary
    sizes = [1_000_003, 70_001, max_n, 2048, 2049, 1]
    for k, n in enumerate(sizes):
        x = _case(n, 100 + k, dev)
        ws_bytes = int(lib.sc_scan_workspace_bytes(n))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        ref = torch.full((n + 1,), -1, dtype=torch.int64, device=dev)
        out = torch.full((n + 1,), -2, dtype=torch.int64, device=dev)
        check(lib.sc_exclusive_scan_i32(n, x.data_ptr(), ref.data_ptr(), ref.data_ptr() + 8 * n, ws.data_ptr(),
                                        ws_bytes, st), "sc_exclusive_scan_i32")
        torch.cuda.synchronize()
        alloc0 = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
        check(lib.sc_exclusive_scan_i32_noalloc(n, x.data_ptr(), out.data_ptr(), out.data_ptr() + 8 * n, st),
              "sc_exclusive_scan_i32_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == alloc0
        assert torch.equal(out, ref), n
        # and against the definition
        want = torch.cumsum(x.to(torch.int64), 0)
        assert torch.equal(out[1:], want), n
        assert int(out[0]) == 0

    # n == 0 writes only the total
    tot = torch.full((1,), 7, dtype=torch.int64, device=dev)
    check(lib.sc_exclusive_scan_i32_noalloc(0, None, None, tot.data_ptr(), st), "sc_exclusive_scan_i32_noalloc")
    torch.cuda.synchronize()
    assert int(tot[0]) == 0


def test_noalloc_scan_rejects_sizes_beyond_the_builtin_scratch():
    from street_crafter_b200._lib import lib

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    st = torch.cuda.current_stream().cuda_stream
    max_n = int(lib.sc_exclusive_scan_i32_noalloc_max_n())
    x = torch.zeros(16, dtype=torch.int32, device=dev)     # never read: the size check comes first
    out = torch.zeros(17, dtype=torch.int64, device=dev)
    rc = lib.sc_exclusive_scan_i32_noalloc(max_n + 1, x.data_ptr(), out.data_ptr(), None, st)
    assert rc == -3  # SC_ERR_WORKSPACE
    msg = lib.sc_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(max_n) in msg
    assert lib.sc_exclusive_scan_i32_noalloc(-1, x.data_ptr(), out.data_ptr(), None, st) == -1  # SC_ERR_BAD_ARG
    torch.cuda.synchronize()
