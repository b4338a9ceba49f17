"""The driver's entry point must stay runnable: smoke() is exercised with the rest of the GPU suite."""
import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@pytest.mark.gpu
def test_graft_entry_smoke():
    import __graft_entry__
    __graft_entry__.smoke()
