"""vgan_b200_subspace_dist_static: the subspace distribution without a scratch argument and without any allocation
(scratch = a static device array of the library, include/vgan_b200.h).  No counterpart in the reference; the existing
entry point vgan_b200_subspace_dist (caller scratch) is the yardstick, np.unique the second one.  Integer work: the
outputs must be identical.  The file sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _masks(n, D, p_on, seed):
    rs = np.random.RandomState(seed)
    base = rs.rand(max(n // 4, 1), D) < p_on                       # few distinct rows -> many duplicates
    masks = base[rs.randint(0, base.shape[0], size=n)]
    masks[: n // 5] = rs.rand(n // 5, D) < p_on                    # plus some unique ones
    return masks


class _Call:
    """Inputs and outputs of one call, allocated up front: nothing is allocated between `before` and `after`."""

    def __init__(self, masks):
        from vgan_b200 import _lib
        self.lib, self.L = _lib, _lib.lib()
        self.n, self.D = masks.shape
        self.src = torch.tensor(masks.astype(np.uint8), device="cuda")
        nbytes = C.c_size_t()
        _lib.check(self.L.vgan_b200_subspace_scratch_bytes(self.n, self.D, C.byref(nbytes)), "subspace_scratch_bytes")
        self.scratch_bytes = nbytes.value
        self.scratch = torch.empty(nbytes.value + 16, dtype=torch.uint8, device="cuda")
        self.out = [self._outputs() for _ in range(2)]              # [existing entry point, static one]
        self.ctx = _lib.ctx(torch.cuda.current_device())
        self.stream = torch.cuda.current_stream().cuda_stream

    def _outputs(self):
        # poisoned, so that a row or count the kernels did not write cannot pass by accident
        return (torch.full((self.n, self.D), 7, dtype=torch.uint8, device="cuda"),
                torch.full((self.n,), -1, dtype=torch.int64, device="cuda"),
                torch.full((1,), -1, dtype=torch.int32, device="cuda"))

    def run_existing(self):
        u, c, m = self.out[0]
        self.lib.check(self.L.vgan_b200_subspace_dist(self.ctx, self.src.data_ptr(), 1, self.n, self.D, 0.0,
                                                      self.scratch.data_ptr(), u.data_ptr(), c.data_ptr(),
                                                      m.data_ptr(), self.stream), "subspace_dist")

    def run_static(self):
        u, c, m = self.out[1]
        return self.L.vgan_b200_subspace_dist_static(self.ctx, self.src.data_ptr(), 1, self.n, self.D, 0.0,
                                                     u.data_ptr(), c.data_ptr(), m.data_ptr(), self.stream)

    def results(self, which):
        u, c, m = self.out[which]
        k = int(m.item())
        return u[:k].cpu().numpy().astype(bool), c[:k].cpu().numpy()


def test_static_scratch_entry_point_matches_existing_twice_in_a_row():
    """Two calls in a row through the C ABI, the second on a smaller, different problem (the shared array still
    holds the first call's packed rows and counts): both must equal vgan_b200_subspace_dist and np.unique."""
    from vgan_b200 import _lib
    calls = [_Call(_masks(1000, 130, 0.3, seed=11)), _Call(_masks(500, 20, 0.5, seed=12))]
    assert all(c.scratch_bytes <= _lib.SUBSPACE_STATIC_SCRATCH_BYTES for c in calls)
    masks = [c.src.cpu().numpy().astype(bool) for c in calls]
    for c in calls:
        c.run_existing()
    torch.cuda.synchronize()
    free_before = None
    for k, c in enumerate(calls):
        if k == 1:                                                  # module loaded, symbol address known by now
            free_before = torch.cuda.mem_get_info()[0]
        _lib.check(c.run_static(), "subspace_dist_static")
        torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "the static entry point took device memory"
    for c, m in zip(calls, masks):
        rows0, counts0 = c.results(0)
        rows1, counts1 = c.results(1)
        assert rows1.shape == rows0.shape and np.array_equal(rows1, rows0) and np.array_equal(counts1, counts0)
        want_rows, want_counts = np.unique(m, axis=0, return_counts=True)
        assert np.array_equal(rows1, want_rows) and np.array_equal(counts1, want_counts)
    # same problem again, straight after the other one: identical output once more
    first = calls[0]
    for t in first.out[1]:
        t.fill_(3)
    _lib.check(first.run_static(), "subspace_dist_static")
    rows1, counts1 = first.results(1)
    rows0, counts0 = first.results(0)
    assert np.array_equal(rows1, rows0) and np.array_equal(counts1, counts0)


def test_static_scratch_entry_point_rejects_sizes_beyond_the_array():
    """n = 4096 fits up to D = 1984 (31 packed words + 8 bytes per row = 1 MiB); one more word per row does not."""
    from vgan_b200 import _lib
    fits = _Call(np.zeros((4096, 1984), dtype=bool))
    assert fits.scratch_bytes == _lib.SUBSPACE_STATIC_SCRATCH_BYTES
    _lib.check(fits.run_static(), "subspace_dist_static")
    rows, counts = fits.results(1)
    assert rows.shape == (1, 1984) and not rows.any() and counts.tolist() == [4096]
    too_big = _Call(np.zeros((4096, 1985), dtype=bool))
    assert too_big.scratch_bytes > _lib.SUBSPACE_STATIC_SCRATCH_BYTES
    assert too_big.run_static() == -1                               # VGAN_B200_ERR_ARG
    msg = _lib.lib().vgan_b200_last_error().decode()
    assert "static array holds" in msg and "1048576" in msg
    assert int(too_big.out[1][2].item()) == -1                      # nothing was launched
    too_big.run_existing()                                          # the caller-scratch entry point still takes it
    assert int(too_big.out[0][2].item()) == 1
