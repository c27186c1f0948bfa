"""`tadataka.vo.semi_dense.fusion` as the reference's examples import it
(examples/semi_dense_depth.py:8,41-42); the module itself is absent from the reference checkout
(SURVEY F2, Appendix B).  Array form: fusion(mu1, mu2, var1, var2) -> (mu, var), the argument order
of `fusion_arrays` (src/semi_dense/fusion.rs:13-42)."""
from tadataka_b200.semi_dense import fusion  # noqa: F401
