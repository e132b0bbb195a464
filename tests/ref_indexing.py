"""The independent restatement of the reference's index maps lives in oracle/indexing.py (the reference arm of bench.py
builds its problem from it too); re-exported here for the tests that grew up with this name."""
from oracle.indexing import *  # noqa: F401,F403
from oracle.indexing import _fn, _spline_rows  # noqa: F401
