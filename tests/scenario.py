"""The API harness lives in the package (navigation_b200/harness.py) so that
bench.py and __graft_entry__.smoke() do not import from tests/; the tests keep
using it under this name."""
from navigation_b200.harness import LASER_DEFAULTS, Side, make_scan, match_clusters  # noqa: F401
