"""CPU: the C restatement of SwAligner::ungappedAlign (oracle/dp_oracle.c) against dumps of the unmodified reference
(tests/golden/ungapped_{ee,loc}.txt, oracle/make_ungapped_goldens.py)."""
import os

import pytest

import helpers as H
from bt2gpu import scoring


@pytest.mark.parametrize("tag,local", [("ee", False), ("loc", True)])
def test_oracle_ungapped_matches_reference(tag, local):
    sc = scoring.Scoring(local=local)
    seq, qual, win, jobs, exp = H.parse_ungapped_golden(os.path.join(H.GOLDEN, "ungapped_%s.txt" % tag), sc)
    assert len(jobs) == len(exp) == 600
    res, edits = H.oracle_ungapped(sc.to_struct(), seq, qual, win, jobs)
    for i in range(len(jobs)):
        assert H.ungapped_to_py(res, edits, i) == exp[i], (i, jobs[i])
    assert {e[0] for e in exp} >= ({1, 0, -1} if local else {1, 0})
