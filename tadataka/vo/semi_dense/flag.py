"""`tadataka.vo.semi_dense.flag` for installations without the reference's Python package.
(Namespace-package layout on purpose: no __init__.py here, so a real `tadataka` install wins.)"""
from tadataka_b200.flag import ResultFlag  # noqa: F401
