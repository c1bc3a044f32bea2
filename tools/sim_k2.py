"""CPU simulation of the K2 folding recurrence: checks that the 64-word LFSR modulo
Q' = y^64 + y^48 + y^5 + x^2 (y^54 + y^37 + y^26)   (= y^5 * Q, Q found by tools/find_word_aligned_multiples.c)
leaves a state congruent to the message modulo the CRC-32C generator."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle

M, R = 64, 2
ALIGNED = [48, 5]
SHIFTED = [54, 37, 26]
MASK = 0xFFFFFFFF


def funnel_r(lo, hi, r):
    return ((lo >> r) | (hi << (32 - r))) & MASK


def fold(words):
    s = [0] * M          # s[k] = coefficient word of y^k, reflected bit order (bit 31 = x^0 ... as loaded LE)
    t_prev = 0
    for w in words:
        t = s[M - 1]
        s = [int(w)] + s[:-1]
        for j in ALIGNED:
            s[j] ^= t
        f = funnel_r(t_prev, t, R)
        for i in SHIFTED:
            s[i + 1] ^= f
        t_prev = t
    for i in SHIFTED:    # pending low part of the last top word
        s[i] ^= t_prev >> R
    return s


def raw_of_words(words_msb_first):
    data = np.array(words_msb_first, dtype=np.uint32).tobytes()
    return (~oracle.crc32c(data, 0xFFFFFFFF)) & MASK   # register after the bytes, starting from 0


rng = np.random.default_rng(1)
for nwords in (1, 5, 63, 64, 65, 128, 1000, 4096):
    w = rng.integers(0, 2**32, size=nwords, dtype=np.uint64).astype(np.uint32)
    s = fold(w)
    got = raw_of_words([s[k] for k in range(M - 1, -1, -1)])
    want = raw_of_words(list(w))
    print(nwords, hex(got), hex(want), "OK" if got == want else "MISMATCH")
    assert got == want
print("K2 recurrence verified")
