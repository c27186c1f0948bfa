"""`rust_bindings.interpolation` (reference: src/py/interpolation.rs:18-22)."""
from tadataka_b200.helpers import interpolation  # noqa: F401
