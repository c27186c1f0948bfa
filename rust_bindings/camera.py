"""`rust_bindings.camera` (reference: src/py/camera.rs:37-41)."""
from tadataka_b200.camera import CameraParameters  # noqa: F401
