"""Host-side model of the producer / issuer hand-shake of ``csrc/stem.cu`` (the stem weight-gradient kernel).

The kernel keeps dy row *pairs* in a ring of PAIR_SLOTS slots and xs rows in a ring of NB stages; the TMA producer
runs PAIR_LOOKAHEAD rows ahead and blocks on slots the MMA issuer has not released yet.  A bookkeeping slip there is
not a wrong number but a hung GPU, so the schedule is replayed here for many (batch, size, grid) combinations:
no deadlock, every pair the issuer touches is the one the producer put into that slot, every (dy row, xs row) product
of the convolution is issued exactly once and lands in the accumulator half the reduction kernel reads for it.
This mirrors the C++ logic (``PairSeq``, ``row_uses``, the loops of the two roles); it does not execute it.
"""
import itertools

import pytest

PAIR_SLOTS = 5
PAIR_LOOKAHEAD = 2
ACC_A, ACC_B, ACC_C, ACC_D, ACC_E = range(5)
# (accumulator, half) -> kernel row bh, as read by k_stem_wgrad_reduce
BH_OF = {(ACC_A, 1): 0, (ACC_D, 0): 0, (ACC_A, 0): 1, (ACC_C, 1): 1, (ACC_B, 1): 2, (ACC_C, 0): 2, (ACC_B, 0): 3, (ACC_E, 1): 3}


def stem_grid(n, oh):
    return max(1, min(148, n * (oh + 3) // 32))


def start(r0, J):
    j0 = r0 >> 1
    lo = j0 - 1 if r0 & 1 else j0 - 2
    hi = j0 + 1 if r0 & 1 else j0
    lo = max(lo, 0)
    hi = min(hi, J)
    hi = max(hi, lo)
    return lo, hi


def introduces(r, J):
    return not (r & 1) and (r >> 1) < J


def uses(r, J):
    j, odd = r >> 1, r & 1
    out = []
    for k in range(3):
        pj = j - k
        if 0 <= pj < J and not (k == 2 and odd):
            out.append((pj, [ACC_A if odd else ACC_D, ACC_B if odd else ACC_C, ACC_E][k]))
    return out


def producer(R0, R1, HB, J, NB, st):
    """Generator: yields ('wait', predicate) until the predicate holds, then performs the load."""
    n, r = divmod(R0, HB)
    q = 0

    def load_pair(img, j):
        nonlocal q
        s = q % PAIR_SLOTS
        want = q // PAIR_SLOTS                      # slot must have been released this many times
        yield lambda s=s, want=want: st["pair_released"][s] >= want
        st["pair_slot"][s] = (img, j)
        st["pair_loaded"][s] += 1
        q += 1
    if R0 < R1:
        lo, hi = start(r, J)
        for j in range(lo, hi):
            yield from load_pair(n, j)
    cn, cr, cR = n, r, R0
    for i, R in enumerate(range(R0, R1)):
        ahead = min(R + PAIR_LOOKAHEAD, R1 - 1)
        while cR <= ahead:
            if introduces(cr, J):
                yield from load_pair(cn, cr >> 1)
            cr += 1
            if cr == HB:
                cr, cn = 0, cn + 1
            cR += 1
        s = i % NB
        yield lambda s=s, want=i // NB: st["b_released"][s] >= want
        st["b_slot"][s] = R
        st["b_loaded"][s] += 1


def issuer(R0, R1, HB, J, NB, st, products):
    n, r = divmod(R0, HB)
    q0 = q_next = 0
    if R0 < R1:
        lo, hi = start(r, J)
        q0, q_next = -lo, hi - lo
    for i, R in enumerate(range(R0, R1)):
        if R != R0 and r == 0:
            q0 = q_next
        if introduces(r, J):
            q_next += 1
        s = i % NB
        yield lambda s=s, want=i // NB + 1: st["b_loaded"][s] >= want
        assert st["b_slot"][s] == R
        for pj, acc in uses(r, J):
            q = q0 + pj
            assert q >= 0
            ps = q % PAIR_SLOTS
            yield lambda ps=ps, want=q // PAIR_SLOTS + 1: st["pair_loaded"][ps] >= want
            assert st["pair_slot"][ps] == (n, pj), (st["pair_slot"][ps], (n, pj), r)
            for half in (0, 1):                       # dy row 2 pj + half times xs row r
                oh = 2 * pj + half
                products.append((n, oh, r, acc, half))
        st["b_released"][s] += 1
        if not (r & 1):
            jl = (r >> 1) - 2
            if 0 <= jl < J:
                st["pair_released"][(q0 + jl) % PAIR_SLOTS] += 1
        r += 1
        if r == HB:
            r, n = 0, n + 1


def run_cta(R0, R1, HB, J, NB):
    st = {"pair_released": [0] * PAIR_SLOTS, "pair_loaded": [0] * PAIR_SLOTS, "pair_slot": [None] * PAIR_SLOTS,
          "b_released": [0] * NB, "b_loaded": [0] * NB, "b_slot": [None] * NB}
    products = []
    roles = [producer(R0, R1, HB, J, NB, st), issuer(R0, R1, HB, J, NB, st, products)]
    waiting = [None, None]
    done = [False, False]
    while not all(done):
        progressed = False
        for k, role in enumerate(roles):
            if done[k]:
                continue
            while True:
                if waiting[k] is not None:
                    if not waiting[k]():
                        break
                    waiting[k] = None
                    progressed = True
                try:
                    waiting[k] = next(role)
                except StopIteration:
                    done[k] = True
                    progressed = True
                    break
        assert progressed, f"deadlock: rows [{R0}, {R1}) HB={HB} NB={NB} state={st}"
    return products


def check(n_img, oh, grid, nb):
    HB, J = oh + 3, oh // 2
    total = n_img * HB
    seen = {}
    for b in range(grid):
        R0, R1 = total * b // grid, total * (b + 1) // grid
        for (n, dy_row, r, acc, half) in run_cta(R0, R1, HB, J, nb):
            bh = r - dy_row
            if (acc, half) in BH_OF:
                assert BH_OF[(acc, half)] == bh, (acc, half, bh)
                key = (n, dy_row, bh)
                assert key not in seen, key
                seen[key] = True
            else:
                assert bh in (-1, 4), (acc, half, bh)          # the two unused accumulator halves
    # every (image, dy row, kernel row) product of the convolution exactly once
    assert len(seen) == n_img * oh * 4


@pytest.mark.parametrize("oh", [2, 4, 16, 32, 48, 64, 112, 144])
def test_schedule_is_complete_and_deadlock_free(oh):
    for n_img in (1, 2, 3, 7, 32):
        grids = {stem_grid(n_img, oh), 1, 2, 5, min(148, n_img * (oh + 3))}
        for grid, nb in itertools.product(sorted(grids), (2, 5, 8)):
            check(n_img, oh, grid, nb)


def test_production_shapes():
    for n_img, oh, nb in ((224, 112, 5), (512, 64, 8), (128, 144, 2), (256, 112, 5)):
        check(n_img, oh, stem_grid(n_img, oh), nb)
