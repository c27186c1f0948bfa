"""`rust_bindings.warp` (reference: src/py/warp.rs:69-75)."""
from tadataka_b200.helpers import warp_vec, warp_vecs  # noqa: F401
