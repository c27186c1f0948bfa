"""`rust_bindings.semi_dense` (reference: src/py/semi_dense.rs:235-246)."""
from tadataka_b200.semi_dense import (Frame, Params, estimate_debug_, fusion,  # noqa: F401
                                      increment_age, propagate, regularize, update_depth)
