"""The geometry the scene-mode sweep relies on (csrc/prox_dev.cuh: scl_octant, scl_wide_range, scl_k_build_wlev), checked
on the host against the weight table built with the reference's formula (operator.py:540-621 via
_native.monotonic_tables_host): (1) an offset's non-zero weights sit on the four candidate taps of its octant -- the
three one step back along the long axis and the one a step back along the short axis; (2) the level / octant / a
enumeration with the duplicate images on the octant boundaries dropped visits every offset but the peak exactly
once, in an order in which every strictly closer neighbour comes at a lower level.  No GPU."""
import numpy as np
import pytest

from scarlet_b200 import _native as nat

COORDS = nat._COORDS


def octant(o):
    sb, sa, sw = o & 1, (o >> 1) & 1, (o >> 2) & 1
    ya, xa = (0, -1 if sa else 1) if sw else (-1 if sa else 1, 0)
    yb, xb = (-1 if sb else 1, 0) if sw else (0, -1 if sb else 1)
    taps = []
    for da, db in ((-1, -1), (-1, 0), (-1, 1), (0, -1)):
        oy, ox = da * ya + db * yb, da * xa + db * xb
        taps.append(COORDS.index((oy, ox)))
    return sb, sw, (ya, xa, yb, xb), sorted(taps)


def level_range(v, sb, sw):
    lo = (v + 2) // 3 + (1 if (sw and v % 3 == 0) else 0)
    hi = v // 2 - (1 if (sb and v % 2 == 0) else 0)
    return lo, hi


@pytest.mark.parametrize("shape", [(9, 9), (12, 17), (33, 20)])
def test_four_candidate_taps_and_level_order(shape):
    Ny, Nx = shape
    wt9, _ = nat.monotonic_tables_host(Ny, Nx)
    amax = max(Ny, Nx) - 1
    seen = np.zeros((2 * Ny - 1, 2 * Nx - 1), dtype=int)
    level = np.full((2 * Ny - 1, 2 * Nx - 1), -1)
    for o in range(8):
        sb, sw, (ya, xa, yb, xb), taps = octant(o)
        others = [i for i in range(8) if i not in taps]
        for v in range(1, 3 * amax + 1):
            lo, hi = level_range(v, sb, sw)
            for a in range(lo, min(hi, amax) + 1):
                bb = v - 2 * a
                Yo, Xo = a * ya + bb * yb, a * xa + bb * xb
                if abs(Yo) > Ny - 1 or abs(Xo) > Nx - 1:
                    continue
                seen[Yo + Ny - 1, Xo + Nx - 1] += 1
                level[Yo + Ny - 1, Xo + Nx - 1] = v
                for bt in range(9):             # interior and every border type
                    assert not wt9[bt, Yo + Ny - 1, Xo + Nx - 1, others].any()
    seen[Ny - 1, Nx - 1] += 1                    # the peak itself is not swept
    assert (seen == 1).all()
    # every tap that carries weight points to a strictly lower level (1 to 3 below)
    level[Ny - 1, Nx - 1] = 0
    ys, xs, ts = np.nonzero(wt9[0])
    for y, x, t in zip(ys, xs, ts):
        oy, ox = COORDS[t]
        if not (0 <= y + oy < 2 * Ny - 1 and 0 <= x + ox < 2 * Nx - 1):
            continue        # an offset no frame of this shape holds (its pixel would be on the frame edge: border tables)
        d = level[y, x] - level[y + oy, x + ox]
        assert 1 <= d <= 3
