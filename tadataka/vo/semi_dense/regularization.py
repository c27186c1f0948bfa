"""`tadataka.vo.semi_dense.regularization` (imported by examples/semi_dense_vo.py:19; absent from
the reference checkout).  regularize(depth_map, variance_map, flag_map) -> depth_map,
the signature of the reference's commented-out binding (src/py/semi_dense.rs:220-233)."""
from tadataka_b200.semi_dense import regularize  # noqa: F401
