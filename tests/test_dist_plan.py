"""Host-side logic of the distributed theta-only chain (config-4 sizes), checked on the CPU through the C-ABI test hook
gpirt_debug_dist_plan: the shares of the ranks are pure functions of (n, world, rank) — every rank derives the same partition without talking to
anyone — so they must tile the work exactly: every grid-point column, every Cholesky block and every column of every split inversion level once."""
import ctypes as C

import pytest

from gpirt_b200 import _lib

NGRID = 1001


def _plan(n, world, rank):
    out = (C.c_int32 * 64)()
    k = _lib.lib().gpirt_debug_dist_plan(n, world, rank, out, 64)
    assert k >= 4, _lib.lib().gpirt_last_error()
    levels = [(out[4 + 3 * l], out[5 + 3 * l], out[6 + 3 * l]) for l in range(out[3])]
    assert k == 4 + 3 * len(levels)
    return dict(kstar=(out[0], out[1]), chol_blocks=out[2], levels=levels)


@pytest.mark.parametrize("n", [4096, 4203, 10000, 20000])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_ranks_tile_the_distributed_chain_exactly(n, world):
    plans = [_plan(n, world, r) for r in range(world)]
    # grid-point columns of L^-1 k* / K^-1 k*: consecutive runs of whole 64-column tiles
    lo = 0
    for p in plans:
        a, b = p["kstar"]
        assert a == min(lo, NGRID) and a <= b <= NGRID and (a % 64 == 0) and (b % 64 == 0 or b == NGRID)
        lo = b
    assert lo == NGRID
    # 256-column outer blocks of the Cholesky: owned cyclically, all of them once
    blocks = (n + 255) // 256
    assert sum(p["chol_blocks"] for p in plans) == blocks
    assert max(p["chol_blocks"] for p in plans) - min(p["chol_blocks"] for p in plans) <= 1
    # split inversion levels: the same levels on every rank, each level's b columns cut into consecutive ranges at tile boundaries
    widths = [tuple(b for b, _, _ in p["levels"]) for p in plans]
    assert len(set(widths)) == 1
    expected = [b for b in (64 << k for k in range(20)) if b < n and world > 1 and b >= 512 and b // 64 >= world]
    assert list(widths[0]) == expected
    for l, b in enumerate(expected):
        at = 0
        for p in plans:
            _, c_lo, c_hi = p["levels"][l]
            assert c_lo == at and c_lo <= c_hi <= b and c_lo % 64 == 0 and c_hi % 64 == 0
            at = c_hi
        assert at == b
        if b // 64 >= 4 * world:  # enough tiles to balance: 3x - x^2 work model, no rank more than 1.5x the mean
            work = [(3 * (c_hi / b) - (c_hi / b) ** 2) - (3 * (c_lo / b) - (c_lo / b) ** 2) for _, c_lo, c_hi in (p["levels"][l] for p in plans)]
            assert max(work) <= 1.5 * (2.0 / world)


def test_dist_plan_rejects_bad_arguments():
    out = (C.c_int32 * 64)()
    L = _lib.lib()
    assert L.gpirt_debug_dist_plan(4096, 2, 2, out, 64) == _lib.ERR_INVALID
    assert L.gpirt_debug_dist_plan(0, 2, 0, out, 64) == _lib.ERR_INVALID
    assert L.gpirt_debug_dist_plan(10000, 8, 0, out, 5) == _lib.ERR_INVALID  # cap too small for the split levels
