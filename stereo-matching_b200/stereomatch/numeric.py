"""Small integer helpers (API of reference stereomatch/numeric.py:5-26)."""


def is_power_of_two(num: int) -> bool:
    """True for 1, 2, 4, 8, ... (False for 0)."""
    return num > 0 and (num & (num - 1)) == 0


def next_power_of_2(n: int) -> int:
    """Smallest power of two >= n (1 for n == 0), as the reference's table
    (reference tests/test_numeric.py:11-17) expects."""
    if n <= 0:
        return 1
    return 1 << (n - 1).bit_length()
