"""CPU: the kernel LOGIC (planner + fill lanes + traceback, the same SW_HD code the
CUDA kernels compile) run through the lock-step warp emulator in tests/emu and
compared with the reference-generated golden vectors and with the oracle.
The emulator is test infrastructure; the product has no CPU path."""
import random
from collections import defaultdict

import pytest

import emu_util as E


def _by_params(golden):
    by = defaultdict(list)
    for c in golden["cases"]:
        by[tuple(c["params"])].append(c)
    return by


@pytest.mark.parametrize("allow_s16,clean_shortcut,through_state", [(True, True, False), (False, True, False), (True, False, False), (True, False, True)])
def test_emulated_kernels_match_golden(golden, allow_s16, clean_shortcut, through_state):
    """clean_shortcut: paths without a gap are written down from the start cell alone (sw_trace.h clean_path, what
    sw_start_kernel does on the device); off: every pair is walked through hops and block replays; through_state: the
    walk is saved and resumed between two rebuilt blocks, as the device's round kernels hand it from launch to launch"""
    by = _by_params(golden)
    total = clean = 0
    # both arithmetics on the two BASELINE score sets; the 16-bit path additionally on every other score set
    sets = list(by) if allow_s16 else [(10, -15, -30, -5), (200, -150, -260, -11), (3, -1, -2, 0)]
    for p in sets:
        cs = by[p]
        pairs = [(c["ref"].encode("latin1"), c["alt"].encode("latin1")) for c in cs]
        for si, s in enumerate(golden["strategies"]):
            res, st = E.emu_align(pairs, p, s, allow_s16, clean_shortcut=clean_shortcut, through_state=through_state)
            assert (st["s16"] > 0) == allow_s16
            clean += st["clean"]
            bad = [(c["ref"], c["alt"], p, s, r, tuple(c["expect"][si])) for c, r in zip(cs, res) if r != tuple(c["expect"][si])]
            assert not bad, bad[:3]
            total += len(cs)
    assert total > 3000
    assert (clean > 300) if clean_shortcut else (clean == 0)


def test_emulated_mixed_strategies_in_one_batch(golden, oracle):
    cs = [c for c in golden["cases"] if tuple(c["params"]) == (10, -15, -30, -5)][:200]
    pairs = [(c["ref"].encode("latin1"), c["alt"].encode("latin1")) for c in cs]
    strategies = [golden["strategies"][k % 4] for k in range(len(cs))]
    res, _ = E.emu_align(pairs, (10, -15, -30, -5), strategies)
    for k, (c, r) in enumerate(zip(cs, res)):
        assert r == tuple(c["expect"][k % 4])


def test_emulated_strip_mined_long_pairs(oracle):
    """alt lengths around the lane-group and strip boundaries (192 = 8 * 24, 768 = 32 * 24, 1536)."""
    rng = random.Random(11)
    pairs = []
    for ql in (191, 192, 193, 300, 767, 768, 769, 1000, 1537, 2305, 3100):     # from 4 strips on (2305, 3100) the walk hops four blocks at a time
        tl = rng.randint(150, 600)
        ref = bytes(rng.choice(b"ACGT") for _ in range(tl))
        alt = bytearray(rng.choice(b"ACGT") for _ in range(ql))
        w = min(tl, ql)
        alt[:w] = ref[:w]
        for _ in range(6):
            alt[rng.randrange(ql)] = rng.choice(b"ACGT")
        del alt[ql // 3: ql // 3 + 2]
        pairs.append((ref, bytes(alt)))
    for params in ((200, -150, -260, -11), (10, -15, -30, -5)):
        for s in oracle.STRATEGIES:
            res, st = E.emu_align(pairs, params, s)
            for (r, a), got in zip(pairs, res):
                assert got == oracle.align(r, a, params, s), (len(r), len(a), params, s)


def test_emulated_long_gaps_take_the_gap_hops(oracle):
    """deletions and insertions of 7 .. 200 cells starting and ending at every kind of offset against the 8-row / half-lane
    checkpoint grid, in lane groups of 8 and in whole-warp units: the walk crosses whole blocks of such a gap with two
    checkpoint loads (Replay::ehop / fhop) and must come out where the reference's cell-by-cell walk does"""
    from gap_cases import long_gap_pairs
    pairs = long_gap_pairs()
    for params in ((10, -15, -30, -5), (200, -150, -260, -11), (3, -1, -2, 0)):
        for s in oracle.STRATEGIES:
            res, st = E.emu_align(pairs, params, s)
            bad = [(len(r), len(a), params, s, got, oracle.align(r, a, params, s)) for (r, a), got in zip(pairs, res) if got != oracle.align(r, a, params, s)]
            assert not bad, bad[:2]


def test_emulated_long_gaps_across_strips(oracle):
    """the same in strip-mined pairs (alt of 900 .. 2 400 columns: 2 to 4 strips, the last with four-block diagonal hops):
    an insertion hop stays inside its strip, a deletion hop reads the row checkpoints of its own strip"""
    rng = random.Random(9)
    pairs = []
    for n, glen, pos in ((900, 64, 10), (1700, 200, 770), (1700, 33, 1536), (2400, 130, 5), (2400, 300, 1500)):
        core = bytes(rng.choice(b"ACGT") for _ in range(n))
        cut = core[:pos] + core[pos + glen:]
        ins = core[:pos] + bytes(rng.choice(b"ACGT") for _ in range(glen)) + core[pos:]
        short = core[: n // 3]
        pairs += [(cut, core), (core, ins), (short, core), (core[: n // 2], ins)]
    for params in ((200, -150, -260, -11), (10, -15, -30, -5)):
        for s in oracle.STRATEGIES:
            res, st = E.emu_align(pairs, params, s)
            bad = [(len(r), len(a), params, s) for (r, a), got in zip(pairs, res) if got != oracle.align(r, a, params, s)]
            assert not bad, bad[:3]


def test_planner_rejects_bad_input():
    with pytest.raises(RuntimeError):
        E.emu_align([(b"ACGT", b"")], (10, -15, -30, -5), 1)
    with pytest.raises(RuntimeError):
        E.emu_align([(b"ACGT", b"ACGT")], (1 << 27, -(1 << 27), -(1 << 27), -5), 1)


# ---- the device planner (mgl_b200/csrc/sw_plan_dev.h): its stage bodies run on the CPU by the emulator

def test_device_planner_matches_host_planner_structure():
    """Both planners on the same chunk: same launches (arithmetic, lane group, K, pipelined, first unit, units, warps),
    same unit count, same checkpoint / last-row / carry / scratch sizes -- config-2, config-3, mixed and degenerate shapes."""
    rng = random.Random(2024)
    shapes = {
        "config2": [(rng.randint(300, 500), rng.randint(145, 165)) for _ in range(5000)],
        "config3": [(rng.randint(300, 1000), rng.randint(300, 1000)) for _ in range(1500)],
        "mixed": [(rng.randint(300, 500), rng.randint(140, 170)) for _ in range(3000)] + [(rng.randint(300, 1000), rng.randint(300, 1000)) for _ in range(330)]
                 + [(rng.randint(5000, 10000), rng.randint(5000, 10000)) for _ in range(7)] + [(1, 1), (1, 3000), (70000, 5), (5, 70000), (192, 192), (193, 193)],
        "one": [(450, 150)],
        "same": [(400, 150)] * 777,
    }
    for name, lens in shapes.items():
        for params in ((10, -15, -30, -5), (200, -150, -260, -11)):
            for allow in (True, False):
                strategies = [(1, 2, 4, 8)[k % 4] for k in range(len(lens))]
                rng.shuffle(lens)
                rc, host, dev = E.compare_planners(lens, params, strategies, allow)
                assert rc == 0 and host == dev, (name, params, allow, rc, host, dev)


@pytest.mark.parametrize("allow_s16", [True, False])
def test_device_planned_emulated_kernels_match_golden(golden, allow_s16):
    """planner stages + fill lanes + traceback, all the code the GPU runs, against the reference-generated vectors"""
    by = _by_params(golden)
    total = 0
    for p in [(10, -15, -30, -5), (200, -150, -260, -11), (3, -1, -2, 0)]:
        cs = by[p]
        pairs = [(c["ref"].encode("latin1"), c["alt"].encode("latin1")) for c in cs]
        strategies = [golden["strategies"][k % 4] for k in range(len(cs))]
        res, st = E.emu_align(pairs, p, strategies, allow_s16, device_plan=True)
        assert (st["s16"] > 0) == allow_s16
        bad = [(c["ref"], c["alt"], p, r, tuple(c["expect"][k % 4])) for k, (c, r) in enumerate(zip(cs, res)) if r != tuple(c["expect"][k % 4])]
        assert not bad, bad[:3]
        total += len(cs)
    assert total > 1000


def test_device_planned_strip_mined_and_mixed(oracle):
    rng = random.Random(12)
    pairs = []
    for ql in (1, 7, 150, 151, 191, 192, 193, 300, 767, 768, 769, 1000, 1537, 2305):
        for _ in range(2):
            tl = rng.randint(1, 600)
            ref = bytes(rng.choice(b"ACGT") for _ in range(tl))
            alt = bytearray(rng.choice(b"ACGT") for _ in range(ql))
            w = min(tl, ql)
            alt[:w] = ref[:w]
            for _ in range(4):
                alt[rng.randrange(ql)] = rng.choice(b"ACGT")
            pairs.append((ref, bytes(alt)))
    for params in ((200, -150, -260, -11), (10, -15, -30, -5)):
        strategies = [(1, 2, 4, 8)[k % 4] for k in range(len(pairs))]
        res, st = E.emu_align(pairs, params, strategies, device_plan=True)
        for k, ((r, a), got) in enumerate(zip(pairs, res)):
            assert got == oracle.align(r, a, params, strategies[k]), (len(r), len(a), params, strategies[k])


def test_device_planner_rejects_bad_input():
    with pytest.raises(RuntimeError):
        E.emu_align([(b"ACGT", b"ACGT"), (b"ACGT", b"")], (10, -15, -30, -5), 1, device_plan=True)
