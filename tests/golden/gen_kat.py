#!/usr/bin/env python3
"""Generate the known-answer vectors that pin oracle/ (and, through it, the CUDA path).

The reference (mysticrenji/kubernetes-with-nvidia-gpu) holds no golden vector for this
path — its Step 9 pod runs `nvidia-smi` (/root/reference/README.md:313-314) — so these are
authored here, from first principles: every expected value is computed with exact rational
arithmetic (`fractions.Fraction`) and one explicit IEEE-754 round-to-nearest-even step
written out below.  No C, no libm `fmaf`, no numpy rounding is involved, so agreement with
oracle/oracle.c and with the GPU is a real check, not a tautology.

Run:  python tests/golden/gen_kat.py        (rewrites tests/golden/kat_*.json; deterministic)
"""
import json
import os
import random
import struct
from fractions import Fraction

HERE = os.path.dirname(os.path.abspath(__file__))

INF = "inf"


# --------------------------------------------------------------------------- binary32
def f32_decode(bits):
    """bits -> (sign, Fraction) for finite values, (sign, 'inf'), or (0, 'nan')."""
    sign = bits >> 31
    exp = (bits >> 23) & 0xFF
    man = bits & 0x7FFFFF
    if exp == 0xFF:
        return (sign, INF) if man == 0 else (0, "nan")
    if exp == 0:
        mag = Fraction(man, 1 << 149)
    else:
        mag = Fraction((1 << 23) | man, 1) * Fraction(2) ** (exp - 150)
    return sign, mag


def f32_round(sign, mag):
    """Round a non-negative exact magnitude to binary32 bits (RNE), attach sign."""
    if mag == 0:
        return sign << 31
    # find e with 2^e <= mag < 2^(e+1)
    e = mag.numerator.bit_length() - mag.denominator.bit_length()
    if Fraction(2) ** e > mag:
        e -= 1
    if Fraction(2) ** (e + 1) <= mag:
        e += 1
    e = max(e, -126)                      # denormal binade shares the -126 quantum
    quantum = Fraction(2) ** (e - 23)
    q = mag / quantum                     # exact; want nearest integer, ties to even
    n = q.numerator // q.denominator
    rem = q - n
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and (n & 1)):
        n += 1
    if n == (1 << 24):                    # carried into the next binade
        n >>= 1
        e += 1
    if e > 127:
        return (sign << 31) | 0x7F800000
    if n < (1 << 23):                     # denormal (e == -126)
        return (sign << 31) | n
    return (sign << 31) | ((e + 127) << 23) | (n & 0x7FFFFF)


def signed(sign, mag):
    return -mag if sign else mag


def f32_add(a, b):
    sa, ma = f32_decode(a)
    sb, mb = f32_decode(b)
    assert ma != "nan" and mb != "nan"
    if ma == INF or mb == INF:
        if ma == INF and mb == INF:
            assert sa == sb, "inf-inf excluded from the vectors"
            return (sa << 31) | 0x7F800000
        return ((sa if ma == INF else sb) << 31) | 0x7F800000
    x = signed(sa, ma) + signed(sb, mb)
    if x == 0:
        # exact zero: -0 only when both addends are -0 / negative zero sum of like signs
        return (1 << 31) if (sa and sb and ma == 0 and mb == 0) else 0
    return f32_round(1 if x < 0 else 0, abs(x))


def f32_mul(a, b):
    sa, ma = f32_decode(a)
    sb, mb = f32_decode(b)
    assert INF not in (ma, mb) and "nan" not in (ma, mb)
    return f32_round(sa ^ sb, ma * mb)


def f32_fma(s, c, b):
    """round(s*c + b) with ONE rounding."""
    ss, ms = f32_decode(s)
    sc, mc = f32_decode(c)
    sb, mb = f32_decode(b)
    assert INF not in (ms, mc, mb) and "nan" not in (ms, mc, mb)
    psign = ss ^ sc
    x = signed(psign, ms * mc) + signed(sb, mb)
    if x == 0:
        if ms * mc == 0 and mb == 0:
            return (1 << 31) if (psign and sb) else 0
        return 0                           # exact cancellation -> +0 under RNE
    return f32_round(1 if x < 0 else 0, abs(x))


# --------------------------------------------------------------------------- bf16
def bf16_widen(h):
    return h << 16


def bf16_round(bits32):
    """binary32 bits -> bf16 bits, RNE on the dropped 16 bits; NaN -> 0x7FFF."""
    if (bits32 & 0x7FFFFFFF) > 0x7F800000:
        return 0x7FFF
    sign, mag = f32_decode(bits32)
    if mag == INF:
        return (sign << 15) | 0x7F80
    if mag == 0:
        return sign << 15
    # same rounding as f32_round but with an 8-bit significand
    e = mag.numerator.bit_length() - mag.denominator.bit_length()
    if Fraction(2) ** e > mag:
        e -= 1
    if Fraction(2) ** (e + 1) <= mag:
        e += 1
    e = max(e, -126)
    quantum = Fraction(2) ** (e - 7)
    q = mag / quantum
    n = q.numerator // q.denominator
    rem = q - n
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and (n & 1)):
        n += 1
    if n == (1 << 8):
        n >>= 1
        e += 1
    if e > 127:
        return (sign << 15) | 0x7F80
    if n < (1 << 7):
        return (sign << 15) | n
    return (sign << 15) | ((e + 127) << 7) | (n & 0x7F)


# --------------------------------------------------------------------------- generator / checksum (integer spec)
M32 = 0xFFFFFFFF


def mix32(x):
    x ^= x >> 16
    x = (x * 0x7FEB352D) & M32
    x ^= x >> 15
    x = (x * 0x846CA68B) & M32
    x ^= x >> 16
    return x


def stream_key(seed, stream):
    return mix32(seed ^ ((0x9E3779B9 * (stream + 1)) & M32))


def hash32(key, index):
    lo = index & M32
    hi = (index >> 32) & M32
    return mix32(lo ^ key ^ ((hi * 0x9E3779B9) & M32))


def gen_f32(key, index, mode):
    h = hash32(key, index)
    if mode == 0:
        k = h >> 8
        return f32_round(0, Fraction(k, 1 << 24))
    if mode == 2:
        return h & 0x81FFFFFF
    if (h & 0x7F800000) == 0x7F800000:
        h &= ~0x00800000 & M32
    return h


def gen_bf16(key, index, mode):
    h = hash32(key, index)
    if mode == 0:
        k = h >> 24
        return f32_round(0, Fraction(k, 1 << 8)) >> 16
    b = h >> 16
    if mode == 2:
        return b & 0x81FF
    if (b & 0x7F80) == 0x7F80:
        b &= ~0x0080 & 0xFFFF
    return b


def checksum(buf, word_offset):
    n = len(buf)
    pad = (-n) % 4
    data = buf + b"\0" * pad
    s0 = s1 = 0
    for j in range(len(data) // 4):
        (w,) = struct.unpack_from("<I", data, 4 * j)
        s0 += w
        s1 += w * (((word_offset + j) & M32) | 1)
    return s0 & 0xFFFFFFFFFFFFFFFF, s1 & 0xFFFFFFFFFFFFFFFF


# --------------------------------------------------------------------------- case selection
def finite32(rng):
    while True:
        b = rng.getrandbits(32)
        if (b >> 23) & 0xFF != 0xFF:
            return b


def near(rng, lo_exp, hi_exp):
    return (rng.getrandbits(1) << 31) | (rng.randrange(lo_exp, hi_exp + 1) << 23) | rng.getrandbits(23)


HAND_F32 = [
    0x00000000, 0x80000000, 0x00000001, 0x80000001, 0x007FFFFF, 0x00800000, 0x00800001,
    0x3F800000, 0xBF800000, 0x3F800001, 0x3F7FFFFF, 0x40400000, 0x7F7FFFFF, 0xFF7FFFFF,
    0x33800000, 0x34000000, 0x4B000000, 0x4B800000, 0x3EAAAAAB, 0x3DCCCCCD,
]


def build_f32(rng):
    add, scale, triad = [], [], []
    for a in HAND_F32:
        for b in HAND_F32:
            add.append([a, b, f32_add(a, b)])
    for _ in range(300):
        a, b = finite32(rng), finite32(rng)
        add.append([a, b, f32_add(a, b)])
    for _ in range(200):                  # same binade: cancellation and ties
        a = near(rng, 120, 130)
        b = near(rng, 120, 130)
        add.append([a, b, f32_add(a, b)])
    for _ in range(150):                  # denormal neighbourhood
        a, b = near(rng, 0, 2), near(rng, 0, 2)
        add.append([a, b, f32_add(a, b)])
    for _ in range(100):                  # UNIT-mode style operands (k * 2^-24)
        a = f32_round(0, Fraction(rng.getrandbits(24), 1 << 24))
        b = f32_round(0, Fraction(rng.getrandbits(24), 1 << 24))
        add.append([a, b, f32_add(a, b)])

    scalars = [0x40400000, 0x3F800000, 0xC0400000, 0x3EAAAAAB, 0x402DF854, 0x00800000]
    for s in scalars:
        for c in HAND_F32:
            scale.append([s, c, f32_mul(s, c)])
    for _ in range(300):
        s, c = near(rng, 100, 150), near(rng, 60, 190)
        scale.append([s, c, f32_mul(s, c)])
    for _ in range(100):
        s, c = near(rng, 110, 130), near(rng, 0, 20)
        scale.append([s, c, f32_mul(s, c)])

    # triad: keep a guaranteed share of cases where the fused result differs from mul-then-add
    differs = 0
    tries = 0
    while differs < 250 and tries < 200000:
        tries += 1
        s = scalars[tries % 5] if tries % 3 else near(rng, 120, 134)
        c, b = near(rng, 110, 140), near(rng, 110, 140)
        fused = f32_fma(s, c, b)
        unfused = f32_add(f32_mul(s, c), b)
        if fused != unfused:
            triad.append([s, c, b, fused, unfused])
            differs += 1
    for s in scalars[:4]:
        for c in HAND_F32[:14]:
            for b in HAND_F32[:14]:
                triad.append([s, c, b, f32_fma(s, c, b), f32_add(f32_mul(s, c), b)])
    for _ in range(200):
        s, c, b = near(rng, 100, 150), near(rng, 0, 3), near(rng, 0, 3)
        triad.append([s, c, b, f32_fma(s, c, b), f32_add(f32_mul(s, c), b)])
    for _ in range(100):
        s = 0x40400000
        c = f32_round(0, Fraction(rng.getrandbits(24), 1 << 24))
        b = f32_round(0, Fraction(rng.getrandbits(24), 1 << 24))
        triad.append([s, c, b, f32_fma(s, c, b), f32_add(f32_mul(s, c), b)])
    return {
        "comment": "bit patterns (uint32). add:[a,b,a+b] scale:[s,c,s*c] triad:[s,c,b,fma(s,c,b),(s*c)+b]",
        "add": add, "scale": scale, "triad": triad,
    }


def finite16(rng):
    while True:
        b = rng.getrandbits(16)
        if (b >> 7) & 0xFF != 0xFF:
            return b


def build_bf16(rng):
    rne = []
    for base in (0x3F80, 0x3F81, 0x0000, 0x0001, 0x007F, 0x0080, 0x7F7F, 0xBF80, 0x8001):
        for low in (0x0000, 0x0001, 0x7FFF, 0x8000, 0x8001, 0xFFFF):
            b32 = (base << 16) | low
            rne.append([b32, bf16_round(b32)])
    for b32 in (0x7F800000, 0xFF800000, 0x7FC00000, 0x7F800001, 0xFFFFFFFF, 0x7F7FFFFF, 0x7F7F8000, 0x7F7F7FFF):
        rne.append([b32, bf16_round(b32)])
    for _ in range(300):
        b32 = rng.getrandbits(32)
        rne.append([b32, bf16_round(b32)])

    add, scale, triad = [], [], []
    scalars = [0x40400000, 0x3EAAAAAB, 0xC0400000, 0x3F800000]
    for _ in range(400):
        if rng.random() < 0.5:
            a, b = finite16(rng), finite16(rng)
        else:
            a = (rng.getrandbits(1) << 15) | (rng.randrange(120, 131) << 7) | rng.getrandbits(7)
            b = (rng.getrandbits(1) << 15) | (rng.randrange(120, 131) << 7) | rng.getrandbits(7)
        add.append([a, b, bf16_round(f32_add(bf16_widen(a), bf16_widen(b)))])
    for _ in range(100):
        a = (rng.getrandbits(1) << 15) | rng.getrandbits(9)
        b = (rng.getrandbits(1) << 15) | rng.getrandbits(9)
        add.append([a, b, bf16_round(f32_add(bf16_widen(a), bf16_widen(b)))])
    for _ in range(300):
        s = rng.choice(scalars)
        c = (rng.getrandbits(1) << 15) | (rng.randrange(1, 200) << 7) | rng.getrandbits(7)
        scale.append([s, c, bf16_round(f32_mul(s, bf16_widen(c)))])
    for _ in range(400):
        s = rng.choice(scalars)
        c = (rng.getrandbits(1) << 15) | (rng.randrange(110, 140) << 7) | rng.getrandbits(7)
        b = (rng.getrandbits(1) << 15) | (rng.randrange(110, 140) << 7) | rng.getrandbits(7)
        triad.append([s, c, b, bf16_round(f32_fma(s, bf16_widen(c), bf16_widen(b)))])
    return {
        "comment": "rne:[f32 bits, bf16 bits]; add:[a,b,out] scale:[s(f32 bits),c,out] triad:[s(f32 bits),c,b,out]; bf16 as uint16",
        "rne": rne, "add": add, "scale": scale, "triad": triad,
    }


def build_gen():
    out = {"comment": "generator spec: keys, hashes, first values per mode; checksum of byte buffers",
           "keys": [], "hash": [], "f32": [], "bf16": [], "checksum": []}
    for seed, stream in ((0, 0), (1, 0), (1, 1), (20261106, 2), (0xFFFFFFFF, 7)):
        out["keys"].append([seed, stream, stream_key(seed, stream)])
    key = stream_key(20261106, 0)
    idx = [0, 1, 2, 3, 255, 49999, (1 << 28) - 1, (1 << 32) - 1, 1 << 32, (1 << 32) + 1, (1 << 34) + 12345]
    for i in idx:
        out["hash"].append([key, i, hash32(key, i)])
    for mode in (0, 1, 2):
        for i in idx:
            out["f32"].append([key, i, mode, gen_f32(key, i, mode)])
            out["bf16"].append([key, i, mode, gen_bf16(key, i, mode)])
    rng = random.Random(99)
    for n, off in ((0, 0), (1, 0), (3, 5), (4, 0), (7, 1 << 32), (64, (1 << 32) - 8), (1001, 123456789012)):
        buf = bytes(rng.getrandbits(8) for _ in range(n))
        s0, s1 = checksum(buf, off)
        out["checksum"].append([buf.hex(), off, s0, s1])
    return out


def main():
    rng = random.Random(20261106)
    for name, obj in (("kat_f32.json", build_f32(rng)), ("kat_bf16.json", build_bf16(rng)), ("kat_gen.json", build_gen())):
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(obj, f, separators=(",", ":"))
            f.write("\n")
        print(name, {k: len(v) for k, v in obj.items() if isinstance(v, list)})


if __name__ == "__main__":
    main()
