"""Host-side mirror of `pair_images` / `launch_inst` (csrc/sib_kernels_tc.cu): which CTA runs which images of a launch.

The CUDA side is checked on the GPU (tests/test_gpu_tc.py::test_tc_launch_size_does_not_change_an_image); this restates
the arithmetic in Python and proves, for every launch size up to a few rounds, that the runs of images are disjoint,
cover [0, n) exactly, never exceed a pair (16 images) and differ by at most one image between the CTAs of the last
round — the property the change was made for."""
import pytest

K_PAIR = 16


def plan(n: int, sms: int = 148):
    grid = min(n, sms)
    rounds = n // K_PAIR // grid
    rem = n - rounds * grid * K_PAIR
    per = (rem + grid - 1) // grid
    return grid, rounds, per


def pair_images(n, grid, rounds, per, cta, qi):
    if qi < rounds:
        base = (cta + qi * grid) * K_PAIR
        return base, base + K_PAIR
    base = rounds * grid * K_PAIR + cta * per
    lim = min(base + per, n)
    return (base, lim) if (qi == rounds and base < lim) else None


@pytest.mark.parametrize("sms", [148, 132, 1])
def test_every_image_is_run_once(sms):
    sizes = list(range(1, 2600)) + [4735, 4736, 4737, 7109, 65536, 65537, 100000]
    if sms == 1:
        sizes = list(range(1, 200))
    for n in sizes:
        grid, rounds, per = plan(n, sms)
        assert 0 <= per <= K_PAIR and n - rounds * grid * K_PAIR < grid * K_PAIR
        seen = bytearray(n)
        last = []
        for cta in range(grid):
            qi = 0
            while True:
                r = pair_images(n, grid, rounds, per, cta, qi)
                if r is None:
                    break
                b, l = r
                assert 0 <= b < l <= n and l - b <= K_PAIR
                for i in range(b, l):
                    assert not seen[i]
                    seen[i] = 1
                if qi == rounds:
                    last.append(l - b)
                qi += 1
            assert pair_images(n, grid, rounds, per, cta, qi + 1) is None
        assert all(seen), n
        if last:  # the remainder is spread evenly: full runs, then at most one shorter run, then nothing
            assert max(last) == per and sum(1 for x in last if x < per) <= 1


def test_small_launches_use_one_cta_per_image():
    for n in (1, 2, 16, 147, 148):
        grid, rounds, per = plan(n)
        assert (grid, rounds, per) == (n, 0, 1)
    assert plan(149) == (148, 0, 2) and plan(2368) == (148, 1, 0) and plan(65536) == (148, 27, 11)
