"""CPU: the PairHMM oracle (oracle/pairhmm_oracle.c) against the reference's own fixture values, against outputs of
the reference itself (tests/golden/pairhmm_golden.json.gz, made by tests/golden/make_pairhmm_golden.py) and, when
oracle/_ref was built here, against the live reference.  Tolerance 1e-5 on the log10 likelihood: the bar the
reference's own test uses (MicrosoftPairHmmUnitTest.java:101)."""
import gzip
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-5


@pytest.fixture(scope="module")
def ph_golden():
    with gzip.open(os.path.join(ROOT, "tests", "golden", "pairhmm_golden.json.gz"), "rb") as f:
        return json.loads(f.read().decode())


@pytest.fixture(scope="module")
def ph():
    from oracle import pairhmm_oracle
    pairhmm_oracle.build()
    return pairhmm_oracle


def _read(r):
    return tuple(x.encode("latin1") for x in r)


def test_reference_fixture_values(ph, ph_golden):
    """the 104 cases of pairhmm-testdata.txt, float and double mode, at the reference test's own tolerance"""
    assert len(ph_golden["file"]) == 104
    for use_double in (False, True):
        for c in ph_golden["file"]:
            got = ph.likelihoods([_read(c["read"])], [c["hap"].encode("latin1")], use_double)[0][0, 0]
            assert abs(got - c["expected_file"]) <= TOL
            assert abs(got - c["ref_double" if use_double else "ref_float"]) <= TOL
    s = ph_golden["simple"]
    got = ph.likelihoods([_read(s["read"])], [s["hap"].encode()], False)[0][0, 0]
    assert abs(got - s["expected"]) <= TOL and abs(got - s["ref_float"]) <= TOL


def test_reference_generated_batches(ph, ph_golden):
    """every read against every haplotype, incl. the pairs whose float result underflows and is redone in double"""
    n = esc_total = 0
    for b in ph_golden["batch"]:
        reads = [_read(r) for r in b["reads"]]
        haps = [h.encode("latin1") for h in b["haps"]]
        keep = np.array(b["finite"], dtype=bool)
        for use_double, key in ((False, "ref_float"), (True, "ref_double")):
            got, esc = ph.likelihoods(reads, haps, use_double)
            exp = np.array([[v if v is not None else np.nan for v in row] for row in b[key]])
            assert np.all(np.abs(got - exp)[keep] <= TOL), (key, float(np.nanmax(np.abs(got - exp)[keep])))
            if not use_double:
                esc_total += int(esc.sum())
        n += int(keep.sum())
    assert n > 1000 and esc_total > 100


def test_oracle_vs_live_reference(ph):
    if not ph.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here); the golden vectors cover this")
    import random
    rng = random.Random(5)
    haps = [bytes(rng.choice(b"ACGTN") for _ in range(rng.randint(8, 300))) for _ in range(7)]
    reads = []
    for _ in range(40):
        n = rng.randint(1, 200)
        g = rng.choice((10, 12))
        reads.append((bytes(rng.choice(b"ACGTN") for _ in range(n)), bytes(rng.randint(6, 40) for _ in range(n)),
                      bytes(rng.randint(20, 50) for _ in range(n)), bytes(rng.randint(20, 50) for _ in range(n)), bytes([g] * n)))
    for use_double in (False, True):
        got, _ = ph.likelihoods(reads, haps, use_double)
        ref = ph.ref_likelihoods(reads, haps, use_double)
        ok = np.isfinite(ref)
        assert ok.mean() > 0.9 and np.all(np.abs(got - ref)[ok] <= TOL)


@pytest.mark.parametrize("rpl", [1, 2, 3, 4, 5, 6, 8, 105, 106, 107, 108, 110])   # 1xx: half-warp groups of 16 lanes
def test_lane_code_through_warp_emulator(ph, ph_golden, rpl):
    """the PhLane code the CUDA kernel compiles (ph_lane.h), run by a lock-step CPU emulation of a warp with every
    rows-per-lane shape (bands of 32/64/128/256 rows: reads of 1..260 bases cross the band boundary), against the
    reference-generated vectors"""
    import emu_util
    for c in ph_golden["file"][::8]:
        got = emu_util.ph_emu([_read(c["read"])], [c["hap"].encode("latin1")], False, rpl)[0, 0]
        assert abs(got - c["expected_file"]) <= TOL
    for b in ph_golden["batch"][:4]:
        reads = [_read(r) for r in b["reads"]]
        haps = [h.encode("latin1") for h in b["haps"]]
        keep = np.array(b["finite"], dtype=bool)
        for use_double, key in ((False, "ref_float"), (True, "ref_double")):
            got = emu_util.ph_emu(reads, haps, use_double, rpl)
            exp = np.array([[v if v is not None else np.nan for v in row] for row in b[key]])
            assert np.all(np.abs(got - exp)[keep] <= TOL), (rpl, key, float(np.nanmax(np.abs(got - exp)[keep])))
