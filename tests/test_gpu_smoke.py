"""The driver's smoke() hook is part of the GPU suite, so a change that breaks it shows up before the round ends."""
import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@pytest.mark.gpu
def test_smoke_hook_runs(capsys):
    import __graft_entry__ as g
    g.smoke()
    out = capsys.readouterr().out
    assert "smoke ok" in out and "libfedrec_b200.so" in out
