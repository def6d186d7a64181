"""The shared-memory layout of the in-CTA transforms (proton_b200/csrc/em_fft.cuh) rests on two properties of the index
padding that the kernels use without checking: (1) the padded distance between the elements of a butterfly does not
depend on the butterfly (so the offsets are compile-time immediates, `pad_off`), and (2) the 16-byte accesses of a
quarter-warp fall into different bank groups.  This mirrors plan_radix / pad_idx of the header and checks both by brute
force for every transform length the library instantiates (CPU only)."""
import pytest


def lg2(r):
    return {2: 1, 4: 2, 8: 3, 16: 4}[r]


def plan(log2p):                      # em_fft.cuh: plan16 / plan_n / plan_radix
    if log2p in (8, 12, 13):
        n = log2p // 4 + (1 if log2p % 4 else 0)
        return [16 if i < log2p // 4 else {1: 2, 2: 4, 3: 8}[log2p % 4] for i in range(n)]
    n = log2p // 3 + (1 if log2p % 3 else 0)
    return [8 if i < log2p // 3 else (2 if log2p % 3 == 1 else 4) for i in range(n)]


def pad(log2p, p):                    # em_fft.cuh: pad_sh1 / pad_sh2 / pad_sh3 / pad_idx_rt
    pl = plan(log2p)
    sh1 = 4 if pl[-1] == 16 else 3
    sh2 = 0 if log2p in (6, 8) else 6 if log2p == 7 else log2p - lg2(pl[0])
    sh3 = 4 if log2p == 13 else 0
    return p + (p >> sh1) + ((p >> sh2) if sh2 else 0) + ((p >> sh3) if sh3 else 0)


def digit_pos(log2p, k):              # em_fft.cuh: digit_pos
    pos, sh = 0, log2p
    for r in plan(log2p):
        sh -= lg2(r)
        pos |= (k & (r - 1)) << sh
        k >>= lg2(r)
    return pos


def worst_conflict(addrs):            # 16-byte elements: 8 lanes per wavefront, bank group = address mod 8
    worst = 1
    for i in range(0, len(addrs), 8):
        cnt = {}
        for a in set(addrs[i:i + 8]):
            cnt[a % 8] = cnt.get(a % 8, 0) + 1
        worst = max(worst, max(cnt.values()))
    return worst


@pytest.mark.parametrize("log2p", range(6, 14))
def test_padding_identity_and_bank_conflicts(log2p):
    P = 1 << log2p
    pl = plan(log2p)
    assert sum(lg2(r) for r in pl) == log2p
    s = log2p
    for R in pl:
        s -= lg2(R)
        S, NB = 1 << s, P // R
        for b in range(NB):
            j = b & (S - 1)
            base = ((b >> s) << (s + lg2(R))) | j
            for q in range(R):
                assert pad(log2p, base + q * S) == pad(log2p, base) + pad(log2p, q * S), (log2p, R, S, b, q)
        worst = 1
        for q in range(R):
            for b0 in range(0, NB - NB % 32 or NB, 32):
                addrs = []
                for b in range(b0, min(b0 + 32, NB)):
                    j = b & (S - 1)
                    addrs.append(pad(log2p, (((b >> s) << (s + lg2(R))) | j) + q * S))
                worst = max(worst, worst_conflict(addrs))
        assert worst == 1, (log2p, R, S, worst)
    # the digit-reversed scatter of the power-of-two path and the natural-order accesses are conflict-free too
    assert max(worst_conflict([pad(log2p, digit_pos(log2p, k)) for k in range(k0, k0 + 32)]) for k0 in range(0, P, 32)) == 1
    assert max(worst_conflict([pad(log2p, k) for k in range(k0, k0 + 32)]) for k0 in range(0, P, 32)) == 1
    assert pad(log2p, P - 1) + 1 <= ((pad(log2p, P - 1) + 8) & ~7)      # padded_len()


def padded_len(log2p):
    return (pad(log2p, (1 << log2p) - 1) + 8) & ~7


MIX_CLASSES = [(7, 5), (7, 6), (8, 5), (8, 6), (8, 7), (8, 9), (8, 11), (8, 14), (9, 5), (9, 6), (9, 9), (9, 10), (9, 11),
               (9, 12), (9, 14), (9, 16)]                                   # FOR_EACH_MIX of csrc/proton_em.cu


@pytest.mark.parametrize("lr,rows", MIX_CLASSES)
def test_mixed_radix_rows_layout_and_choice(lr, rows):
    """em_mix.cuh: P = rows x 2^lr points as `rows` padded rows.  The column passes read and write element
    (row q, column j) for consecutive j per lane — conflict-free because a row's padding is conflict-free in natural order
    and the row stride is a multiple of 8 elements — and the rows fit the shared memory (two CTAs per SM for the 256-thread classes,
    MixCfg).  The host's choice of class (classify() in
    csrc/proton_em.cu) picks the admissible class of least measured cost."""
    from proton_b200._lib import class_of
    R, rs = 1 << lr, padded_len(lr)
    assert rs % 8 == 0
    for q in range(rows):
        assert max(worst_conflict([q * rs + pad(lr, j) for j in range(j0, j0 + 32)]) for j0 in range(0, R, 32)) == 1
    tw = sum((1 << (lr - sum(lg2(r) for r in plan(lr)[:i + 1]))) for i in range(len(plan(lr)))
             if lr - sum(lg2(r) for r in plan(lr)[:i + 1]) > 0)
    smem = (tw + rows * rs) * 16
    P = rows << lr
    nt = 512 if P > 6144 else 256 if P > 3072 else 128 if P >= 1024 else 64                  # MixCfg::NT / MINB
    minb = (512 if (lr >= 8 or rows >= 9) else 768) // nt
    assert smem + 192 <= 227 * 1024
    if 3072 < P <= 6144:                                       # 256 threads, 128 registers: two CTAs per SM must fit
        assert nt == 256 and minb == 2 and 2 * (smem + 192 + 1024) <= 228 * 1024, (smem, minb)   # (1 KB per CTA is reserved)
    n_max = (P + 1) // 2                                       # 2N - 1 <= P
    n = n_max - 1 if n_max & (n_max - 1) == 0 else n_max       # (a power-of-two line has its own kernel)
    assert class_of(n) == ("fft_mix", lr, rows)                # cheapest admissible class for the longest line it takes


def test_every_line_length_gets_an_admissible_transform():
    """classify() (through proton_em_debug_class_of) for every line length the build supports: the transform of the
    chosen class is long enough for Bluestein (P >= 2N - 1), never longer than the next power of two, and the class
    is one that is built."""
    from proton_b200._lib import class_of
    built = set(MIX_CLASSES)
    for n in list(range(1, 4200)) + list(range(4200, 65537, 37)) + [65535, 65536, 131072]:
        kind, lr, rows = class_of(n)
        pow2 = n & (n - 1) == 0
        need = 2 * n - 1
        if kind == "dense":
            assert n <= 88 and not (pow2 and n >= 64)
        elif kind == "fft_pow2":
            assert pow2 and 64 <= n <= 8192 and (1 << lr) == n
        elif kind == "fft_blue":
            assert not pow2 and (1 << lr) >= need and (1 << (lr - 1)) < need and lr <= 12
        elif kind == "fft_mix":
            P = rows << lr
            assert (lr, rows) in built and need <= P <= 1 << (need - 1).bit_length() and n <= 4096
        elif kind == "long_pow2":
            assert pow2 and n > 8192 and rows * 4096 == n
        else:
            assert kind == "long_blue" and 4096 < n <= 65536 and rows * 4096 >= need and rows <= 32
    assert class_of(65537)[0] == "none" and class_of(262144)[0] == "none"
