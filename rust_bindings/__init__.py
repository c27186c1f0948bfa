"""Import shim: `rust_bindings` as the reference's Python code imports it (setup.py:97-104 builds
eight pyo3 extension modules under this name).  Every submodule re-exports the B200
implementation in `tadataka_b200`; put this repository root on PYTHONPATH ahead of the
reference's compiled extension to switch over.
"""
