"""`rust_bindings.triangulation` (reference: src/py/triangulation.rs:21-26)."""
from tadataka_b200.helpers import calc_depth0  # noqa: F401
