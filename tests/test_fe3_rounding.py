"""The float32 rounding of the integer-limb DFT (honk_b200/csrc/fe_i8.cu: f3_round_f32), restated in Python integers and
checked against exact round-to-nearest-even.  The kernel holds the spectrum as an exact integer T = 256 S + a (S: 64-bit
Horner state, a: the last int32 accumulator) and must produce float32(T * 2^-61) - the complex64 store of librosa 0.6's
stft (manage_audio.py:31-38) applied to the exact value - with one int32 -> float32 conversion and no FP64 instruction."""
import random

import numpy as np


def _s32(x):
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x >> 31 else x


def f3_round_f32(S, a):
    """Line-by-line mirror of the device function (same variable names)."""
    V = S + (_s32(a) >> 8)
    r = a & 255
    vlo = V & 0xFFFFFFFF
    vhi = _s32((V >> 32) & 0xFFFFFFFF)
    fits = vhi == (_s32(vlo) >> 31)
    hi = (_s32(vlo) >> 24) if fits else vhi
    lo = ((vlo << 8) | r) & 0xFFFFFFFF if fits else vlo
    st = 0 if fits else r
    cls = 32 - ((hi ^ (hi >> 31)) & 0xFFFFFFFF).bit_length()
    k = 33 - cls if cls < 32 else ((lo >> 31) ^ (hi & 1))
    lost = (lo << (32 - k)) & 0xFFFFFFFF if k > 0 else 0          # shl.b32 clamps a shift of 32 to zero
    full = ((hi & 0xFFFFFFFF) << 32) | lo
    W = _s32(((full >> k) & 0xFFFFFFFF) | (1 if (lost | st) else 0))
    e = k + (-61 if fits else -53)
    return np.float32(np.float32(W) * np.float32(2.0 ** e))


def _rn_int(z, scale_exp):
    """float32(z * 2^scale_exp) with one round-to-nearest-even from the exact integer z (what I2F.F32.S64 + a power-of-two
    scale give)."""
    if z == 0:
        return np.float32(0)
    sign, m = (-1 if z < 0 else 1), abs(z)
    nb = m.bit_length()
    if nb > 24:
        shift = nb - 24
        q, rem = divmod(m, 1 << shift)
        half = 1 << (shift - 1)
        if rem > half or (rem == half and (q & 1)):
            q += 1
        return np.float32(sign * float(q) * 2.0 ** (shift + scale_exp))
    return np.float32(sign * m * 2.0 ** scale_exp)


def f3_round_f32_xu(S, a):
    """Mirror of the XU-pipe variant (the default): Z = T if |V| < 2^55 else V | sticky, one 64-bit integer conversion."""
    V = S + (_s32(a) >> 8)
    r = a & 255
    fits = -(1 << 55) <= V < (1 << 55)
    Z = (V << 8) | r if fits else (V | (1 if r else 0))
    assert -(1 << 63) <= Z < (1 << 63)
    return _rn_int(Z, -61 if fits else -53)


def exact_rn(S, a):
    T = 256 * S + _s32(a)
    if T == 0:
        return np.float32(0)
    sign, m = (-1 if T < 0 else 1), abs(T)
    nb = m.bit_length()
    if nb <= 24:
        return np.float32(sign * m * 2.0 ** -61)
    shift = nb - 24
    q, rem = divmod(m, 1 << shift)
    half = 1 << (shift - 1)
    if rem > half or (rem == half and (q & 1)):
        q += 1
    return np.float32(sign * float(q) * 2.0 ** (shift - 61))


def test_integer_rounding_matches_exact_rn():
    rnd = random.Random(1)
    for _ in range(120000):
        mag = rnd.choice([0, 1, 5, 10, 20, 22, 23, 24, 25, 30, 31, 32, 33, 40, 50, 60, 62])
        S = rnd.getrandbits(mag) if mag else 0
        if rnd.random() < 0.5:
            S = -S
        if rnd.random() < 0.1:
            S += rnd.choice([-1, 0, 1])
        a = rnd.getrandbits(25)
        if rnd.random() < 0.5:
            a = (-a) & 0xFFFFFFFF
        if rnd.random() < 0.1:
            a = rnd.choice([0, 128, 255, 256, 0xFFFFFF00, 0xFFFFFF80, 1, 0xFFFFFFFF])
        S = max(-(2 ** 62), min(2 ** 62, S))
        assert f3_round_f32(S, a) == exact_rn(S, a), (S, _s32(a))
        assert f3_round_f32_xu(S, a) == exact_rn(S, a), (S, _s32(a))
    # ties: T = (2 q + 1) * 2^(shift - 1) exactly
    for q in (2 ** 23, 2 ** 23 + 1, 2 ** 24 - 1, 2 ** 24 - 2):
        for shift in (1, 2, 7, 8, 9, 16, 30, 40):
            T = (2 * q + 1) << (shift - 1)
            for sgn in (1, -1):
                S, a = divmod(sgn * T, 256)
                assert f3_round_f32(S, a) == exact_rn(S, a), (q, shift, sgn)
                assert f3_round_f32_xu(S, a) == exact_rn(S, a), (q, shift, sgn)
