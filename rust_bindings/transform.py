"""`rust_bindings.transform` (reference: src/py/transform.rs:18-22)."""
from tadataka_b200.helpers import transform  # noqa: F401
