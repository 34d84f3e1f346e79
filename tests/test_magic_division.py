"""The division-free index arithmetic of the kernels (rule 7), restated with Python integers:
common.cuh KMagic (flat thread index -> column, 64-bit multiply-high) and scalars.cu FastDiv (flattened (scalar, level)
index -> level, 32-bit multiply-high) must give the exact quotient on their whole domain."""
import random


def kmagic(nk):
    return (0xFFFFFFFFFFFFFFFF // nk + 1) & 0xFFFFFFFFFFFFFFFF


def umul64hi(a, b):
    return (a * b) >> 64


def test_kmagic_is_exact_for_every_inner_extent_and_any_launch_size():
    rng = random.Random(7)
    for nk in list(range(2, 1100)) + [4096, 65536, 220 * 55]:
        m = kmagic(nk)
        ts = [0, 1, nk - 1, nk, nk + 1, 2 * nk - 1, (1 << 32) - 1, 1 << 32, (1 << 40) + 12345]
        ts += [rng.randrange(0, 1 << 44) for _ in range(50)]
        ts += [c * nk + d for c in (41943042, 125829120, 7864320) for d in (0, nk - 1)]  # column boundaries of the big meshes
        for t in ts:
            assert umul64hi(t, m) == t // nk, (nk, t)


def test_fastdiv_is_exact_for_the_scalar_level_index():
    for ns in range(2, 65):
        magic = ((1 << 32) // ns + 1) & 0xFFFFFFFF
        for x in list(range(0, 5000)) + [ns * 220 + 1, 65535, 1000003]:
            if x * ns < (1 << 32):
                assert (x * magic) >> 32 == x // ns, (ns, x)
