"""CPU tests (no GPU): pin oracle/sc_oracle.c against the reference's golden vectors and, when
oracle/_ref is present, against the reference's own cpu.cu element for element.
Reference behaviour: stream_compaction/stream_compaction/cpu.cu:16-79, harness
stream_compaction/src/main.cpp:22-140 (config 1 of BASELINE.json: n = 2^20 POT and NPOT)."""
import os

import numpy as np
import pytest

from oracle_bindings import ScOracle, gen_array

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "sc_golden.npz"))


@pytest.fixture(scope="module")
def oracle():
    return ScOracle()


def test_generator_matches_golden():
    assert np.array_equal(gen_array(1 << 12, 50), GOLD["scan_in"])
    assert np.array_equal(gen_array(1 << 12, 4), GOLD["compact_in"])
    assert GOLD["scan_in"][-1] == 0          # "Leave a 0 at the end to test that edge case"


def test_scan_golden(oracle):
    a = GOLD["scan_in"]
    assert np.array_equal(oracle.scan(a), GOLD["scan_pot"])
    assert np.array_equal(oracle.scan(a[:-3]), GOLD["scan_npot"])


def test_compact_golden(oracle):
    a = GOLD["compact_in"]
    assert np.array_equal(oracle.compact_without_scan(a), GOLD["compact_pot"])
    assert np.array_equal(oracle.compact_without_scan(a[:-3]), GOLD["compact_npot"])
    assert np.array_equal(oracle.compact_with_scan(a), GOLD["compact_ws_pot"])
    assert np.array_equal(oracle.compact_with_scan(a[:-3]), GOLD["compact_ws_npot"])


def test_wraparound_golden(oracle):
    assert np.array_equal(oracle.scan(GOLD["wrap_in"]), GOLD["wrap_scan"])
    # closed form: exclusive cumsum modulo 2^32
    x = GOLD["wrap_in"].astype(np.int64)
    want = np.concatenate([[0], np.cumsum(x)[:-1]]).astype(np.uint64).astype(np.uint32).view(np.int32)
    assert np.array_equal(GOLD["wrap_scan"], want)


@pytest.mark.parametrize("n", [1, 2, 3, 31, 32, 33, 63, 64, 65])
def test_tiny_golden(oracle, n):
    x = GOLD[f"tiny{n}_in"]
    assert np.array_equal(oracle.scan(x), GOLD[f"tiny{n}_scan"])
    assert np.array_equal(oracle.compact_without_scan(x), GOLD[f"tiny{n}_compact"])
    assert np.array_equal(oracle.compact_with_scan(x), GOLD[f"tiny{n}_compact"])


def test_empty(oracle):
    e = np.zeros(0, dtype=np.int32)
    assert oracle.scan(e).size == 0
    assert oracle.compact_without_scan(e).size == 0
    assert oracle.compact_with_scan(e).size == 0


@pytest.mark.parametrize("n_pot,delta", [(1 << 20, 0), (1 << 20, 3)])
def test_config1_against_reference_build(oracle, n_pot, delta):
    """BASELINE config 1 at full size against the reference's cpu.cu (needs oracle/_ref)."""
    try:
        ref = ScOracle(reference=True)
    except (FileNotFoundError, OSError):
        pytest.skip("oracle/_ref/libref_sc_cpu.so not built here (golden vectors are the pin)")
    a = gen_array(n_pot, 50)[: n_pot - delta]
    assert np.array_equal(oracle.scan(a), ref.scan(a))
    b = gen_array(n_pot, 4)[: n_pot - delta]
    want = ref.compact_without_scan(b)
    assert np.array_equal(oracle.compact_without_scan(b), want)
    assert np.array_equal(oracle.compact_with_scan(b), want)
    assert np.array_equal(ref.compact_with_scan(b), want)
    assert np.array_equal(want, b[b != 0])
