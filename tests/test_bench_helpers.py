"""CPU: the host-side helpers of bench.py that decide what is measured (no GPU, no servers)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def test_sequence_lengths_add_up_and_are_ordered():
    ln = bench.seq_lengths(3100, 24)
    assert len(ln) == 24 and sum(ln) == 3100 * 1000000
    assert all(a >= b for a, b in zip(ln, ln[1:])) and ln[-1] > 20 * 1000000      # a chromosome-like spread, no tiny sequence


def test_clients_never_exceed_the_reference_clients_window():
    """The reference client keeps at most 20000 pairs in flight (pat.h:2464-2550) and its flow control can stall beyond
    that: no client of a bench pass is given more than 16384 pairs (or reads)."""
    for paired in (True, False):
        bench.PAIRED = paired
        for n in (2, 1000, 65536, 262144, 524288, 1048576):
            c = bench.default_clients(n)
            units = n // 2 if paired else n
            assert 2 <= c <= 16
            if units <= 16 * 16384:
                assert -(-units // c) <= 16384, (paired, n, c)
    bench.PAIRED = False


def test_reads_are_reproducible_and_cover_all_sequences():
    bench.PAIRED = True
    try:
        rng = np.random.default_rng(1)
        seqs = [rng.integers(0, 4, size=n, dtype=np.uint8) for n in (60000, 30000, 10000)]
        a, qa = bench.make_reads(5, seqs, 4000)
        b, qb = bench.make_reads(5, seqs, 4000)
        assert a.shape == (4000, bench.RDLEN) and np.array_equal(a, b) and np.array_equal(qa, qb)
        c, _ = bench.make_reads(6, seqs, 4000)
        assert not np.array_equal(a, c)
        assert a.max() <= 4 and qa.min() >= 33
    finally:
        bench.PAIRED = False


def test_workload_name_states_the_configuration():
    bench.PAIRED = True
    try:
        w = bench.workload_name(262144, 3100, 24)
        assert "3100 Mbp" in w and "24 sequences" in w and "131072 pairs" in w and "-X" in w
    finally:
        bench.PAIRED = False
