"""Shared helpers of the parity tests live with the synthetic workload (madpose_b200/workload.py); re-exported here."""
from madpose_b200.workload import models_around_gt, problem, rodrigues  # noqa: F401
