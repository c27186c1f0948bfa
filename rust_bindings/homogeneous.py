"""`rust_bindings.homogeneous` (reference: src/py/homogeneous.rs:36-44)."""
from tadataka_b200.helpers import (from_homogeneous_vec, from_homogeneous_vecs,  # noqa: F401
                                   to_homogeneous_vec, to_homogeneous_vecs)
