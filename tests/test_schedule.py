"""Clock schedule tables (reference pipeline.py:63-79, README.md:72-84)."""
from pipe_b200.runtime.schedule import bubble_fraction, clock_cycles


def test_square():
    assert list(clock_cycles(3, 3)) == [
        [(0, 0)],
        [(1, 0), (0, 1)],
        [(2, 0), (1, 1), (0, 2)],
        [(2, 1), (1, 2)],
        [(2, 2)],
    ]


def test_more_batches_than_stages():
    assert list(clock_cycles(4, 2)) == [
        [(0, 0)], [(1, 0), (0, 1)], [(2, 0), (1, 1)], [(3, 0), (2, 1)], [(3, 1)],
    ]


def test_more_stages_than_batches():
    assert list(clock_cycles(1, 3)) == [[(0, 0)], [(0, 1)], [(0, 2)]]
    assert list(clock_cycles(2, 3)) == [[(0, 0)], [(1, 0), (0, 1)], [(1, 1), (0, 2)], [(1, 2)]]


def test_every_cell_once_and_tick_count():
    for m in range(1, 7):
        for n in range(1, 7):
            ticks = list(clock_cycles(m, n))
            assert len(ticks) == m + n - 1
            cells = [c for t in ticks for c in t]
            assert sorted(cells) == sorted((i, j) for i in range(m) for j in range(n))
            for k, t in enumerate(ticks):
                assert all(i + j == k for i, j in t)
                assert len({j for _, j in t}) == len(t)     # distinct stages in a tick


def test_bubble():
    assert bubble_fraction(4, 2) == 1 / 5
    assert bubble_fraction(7, 8) == 7 / 14
    assert bubble_fraction(1, 1) == 0
