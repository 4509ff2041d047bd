"""The multiplier trick of csrc/conv_igemm.cu (`fast_div`: x / d == (x * M) >> 40 with M = ceil(2^40 / d)) is exact inside the
bounds the host code enforces before it hands out a multiplier (x * d < 2^40 and x / d < 2^23); mirrored here in Python so that a
change of the bounds or of the shift is caught without a GPU."""
import random


def magic(xmax: int, d: int) -> int:
    if d <= 0 or xmax * d >= (1 << 40) or xmax // d >= (1 << 23):
        return 0
    return ((1 << 40) + d - 1) // d


def fast_div(x: int, M: int) -> int:
    prod = x * M
    assert prod < (1 << 64), "the device code multiplies in 64 bits"
    return prod >> 40


def test_fast_div_is_exact_inside_its_bounds():
    rng = random.Random(0)
    checked = 0
    for d in list(range(1, 300)) + [448, 512, 802, 1024, 3136, 12544, 13225]:
        # largest x the host would allow for this divisor (also capped to int32 like the kernels' indices)
        xmax = min((1 << 31) - 1, ((1 << 40) - 1) // d, d * (1 << 23) - 1)
        M = magic(xmax, d)
        if M == 0:
            continue
        xs = {0, 1, d - 1, d, d + 1, xmax, xmax - 1, xmax // 2}
        xs |= {rng.randrange(0, xmax + 1) for _ in range(200)}
        xs |= {k * d + off for k in (1, 7, xmax // d) for off in (-1, 0, 1) if 0 <= k * d + off <= xmax}
        for x in xs:
            assert fast_div(x, M) == x // d, (x, d)
            checked += 1
    assert checked > 10000


def test_magic_refuses_what_it_cannot_guarantee():
    assert magic((1 << 31) - 1, 1 << 12) == 0      # x * d >= 2^40
    assert magic(1 << 24, 1) == 0                   # x / d >= 2^23: the 64-bit product would overflow
    assert magic(3_400_000, 56) != 0                # ResNet-50 at batch 256: positions x image width
