"""`rust_bindings.projection` (reference: src/py/projection.rs:43-51)."""
from tadataka_b200.helpers import (inv_project_vec, inv_project_vecs,  # noqa: F401
                                   project_vec, project_vecs)
