"""CPU: pure host logic of the sharded path added with the peer exchange (hellovulkan_b200/sharded.py):
slab balancing from measured times, chunking of the pipelined light broadcast, phase bookkeeping."""
import numpy as np
import pytest

from hellovulkan_b200.sharded import ShardedCuller, balance_slabs, z_slabs


def covers(slabs, Z):
    assert slabs[0][0] == 0 and slabs[-1][1] == Z
    for (a, b), (c, d) in zip(slabs[:-1], slabs[1:]):
        assert b == c and a <= b
    return True


@pytest.mark.parametrize("Z,world", [(24, 1), (24, 2), (24, 3), (96, 8), (96, 5), (7, 8)])
def test_z_slabs_cover_the_grid(Z, world):
    s = z_slabs(Z, world)
    assert len(s) == world and covers(s, Z)
    sizes = [b - a for a, b in s]
    assert max(sizes) - min(sizes) <= 1


def test_balance_moves_slices_towards_the_cheap_ranks():
    s = z_slabs(96, 2)
    out = balance_slabs(s, [2.0, 3.2])          # the far half is dearer: the near rank takes more slices
    assert covers(out, 96) and out[0][1] > 48
    # piecewise-constant model: costs after the cut are equal within one slice's worth
    dens = [2.0 / 48] * 48 + [3.2 / 48] * 48
    c0 = sum(dens[: out[0][1]]); c1 = sum(dens[out[0][1]:])
    assert abs(c0 - c1) <= max(dens) + 1e-9


def test_balance_is_a_fixed_point_when_balanced_and_keeps_everyone_busy():
    s = z_slabs(96, 8)
    assert balance_slabs(s, [1.0] * 8) == s
    out = balance_slabs(s, [1, 1, 1, 1, 1, 1, 1, 50])      # one very dear slab is cut up, nobody ends up empty
    assert covers(out, 96) and all(b - a >= 1 for a, b in out)
    assert out[-1][1] - out[-1][0] < 12
    assert balance_slabs(s, [0.0] * 8) == s                 # no information: unchanged


def test_balance_converges_on_a_synthetic_cost_profile():
    Z, world = 96, 8
    cost = np.linspace(1.0, 6.0, Z) ** 2                    # far slices much dearer, as in the real grid
    slabs = z_slabs(Z, world)
    worst0 = max(cost[a:b].sum() for a, b in slabs)
    for _ in range(6):
        times = [float(cost[a:b].sum()) for a, b in slabs]
        slabs = balance_slabs(slabs, times)
        assert covers(slabs, Z)
    worst = max(cost[a:b].sum() for a, b in slabs)
    assert worst < 0.6 * worst0 and worst <= 1.25 * cost.sum() / world


@pytest.mark.parametrize("n,chunks", [(0, 4), (1, 4), (1000, 8), (1024, 8), (1025, 2), (5000, 3), (1 << 24, 4), (2048, 64)])
def test_chunk_bounds_are_whole_tiles_in_order(n, chunks):
    b = ShardedCuller.chunk_bounds(n, chunks)
    assert 1 <= len(b) <= max(1, chunks)
    assert b[0][0] == 0 and b[-1][1] == n
    for (a0, a1), (c0, c1) in zip(b[:-1], b[1:]):
        assert a1 == c0 and a0 < a1
    for a0, a1 in b[:-1]:
        assert a0 % 1024 == 0 and a1 % 1024 == 0
    assert b[-1][0] % 1024 == 0


def test_culler_rejects_unknown_modes():
    class Dummy:
        lights = None
    with pytest.raises(ValueError):
        ShardedCuller(Dummy(), (16, 9, 24), world=1, rank=0, exchange="mpi")
    sc = ShardedCuller(Dummy(), (16, 9, 24), world=1, rank=0, slabs=[(0, 24)])
    with pytest.raises(ValueError):
        sc.frame_pipelined(0, 10)                            # needs the peer exchange
    assert sc.phase_ms() == {}


def test_cut_search_converges_on_a_peaked_cost():
    """CutSearch (bench.py's slab balancing): the cost of a frame counted in rounds sits in a few rows; quantiles of the accumulated
    cumulative-share knots find balanced cuts where one piecewise-constant re-cut per iteration keeps overshooting."""
    from hellovulkan_b200.sharded import CutSearch
    rows = 864
    dens = np.full(rows, 0.2)
    dens[560:600] = 30.0            # the edge of the lit region
    dens[600:640] = 6.0
    cum = np.concatenate([[0.0], np.cumsum(dens)])
    world = 8
    cuts = [rows * k // world for k in range(world + 1)]
    search = CutSearch(rows, world)
    worst = []
    for _ in range(10):
        times = [cum[cuts[k + 1]] - cum[cuts[k]] for k in range(world)]
        worst.append(max(times) / (sum(times) / world))
        search.add(cuts, times)
        cuts = search.next_cuts()
        assert cuts[0] == 0 and cuts[-1] == rows and all(b > a for a, b in zip(cuts, cuts[1:]))
    assert worst[0] > 4.0 and min(worst) < 1.25, worst
    assert search.best[0] == min(max(cum[c[k + 1]] - cum[c[k]] for k in range(world)) for c in [search.best[1]])
