"""Pairs with one long gap each -- deletions and insertions of 7 .. 200 cells that start and end at every kind of offset
against the traceback's 8-row / half-lane checkpoint grid, for lane groups of 8 (alt <= 192) and whole-warp units.
TEST INFRASTRUCTURE, shared by the emulator test and the GPU test of the gap hops (Replay::ehop / fhop)."""
import random


def long_gap_pairs(seed=5):
    rng = random.Random(seed)
    pairs = []
    for glen in (7, 8, 9, 15, 16, 17, 31, 33, 64, 65, 130, 200):
        for pos in (1, 2, 7, 8, 9, 40, 77):
            for long_alt in (False, True):
                n = 420 if long_alt else 170
                core = bytes(rng.choice(b"ACGT") for _ in range(n))
                p = min(pos, n - 2)
                cut = core[:p] + core[min(n, p + glen):]                    # the other sequence lacks glen bases at p
                ins = core[:p] + bytes(rng.choice(b"ACGT") for _ in range(glen)) + core[p:]
                if long_alt:
                    pairs += [(cut, core), (core, cut), (ins, core), (core, ins)]
                else:
                    # alt (the second sequence) at most 192 columns: lane groups of 8
                    pairs += [(core + core[:60], cut[:190]), (ins, core[:190]), (cut, core[:190])]
    return [(r, a) for r, a in pairs if len(r) and len(a)]
