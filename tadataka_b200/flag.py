"""Per-pixel outcome codes of the depth update.

Values are those of the reference's `Flag` enum (src/semi_dense/flag.rs:3-14) and of its Python
mirror tadataka/vo/semi_dense/flag.py, member names included (the misspelt SERCH is part of the
reference's API and kept).
"""
import enum


class ResultFlag(enum.IntEnum):
    SUCCESS = 0
    HYPOTHESIS_OUT_OF_SERCH_RANGE = -1
    KEY_OUT_OF_RANGE = -2
    REF_CLOSE_OUT_OF_RANGE = -3
    REF_FAR_OUT_OF_RANGE = -4
    REF_EPIPOLAR_TOO_SHORT = -5
    INSUFFICIENT_GRADIENT = -6
    NEGATIVE_PRIOR_DEPTH = -7
    NEGATIVE_REF_DEPTH = -8
    NOT_PROCESSED = -9

    @classmethod
    def histogram_labels(cls):
        """Labels for the [10] flag histogram the kernels emit (index = -flag)."""
        return [cls(-i).name for i in range(10)]
