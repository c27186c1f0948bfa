"""tadataka_b200 -- Tadataka's semi-dense inverse-depth path, rebuilt for B200 (sm_100a).

    tadataka_b200.semi_dense   drop-in for rust_bindings.semi_dense (numpy in/out)
    tadataka_b200.device       batched, device-resident form (torch CUDA tensors)
    tadataka_b200.helpers      warp / interpolation / projection / transform / triangulation
    tadataka_b200.flag         ResultFlag
    tadataka_b200.scene        synthetic piecewise-planar scenes for tests and benchmarks

All arithmetic runs in libtdk_b200.so (hand-written CUDA behind the C-ABI of include/tdk.h).
"""
__version__ = "0.1.0"
