"""Alias under the reference's (misspelt) module name,
mlcomp/worker/reports/segmenation.py."""
from .segmentation import SegmentationReportBuilder  # noqa: F401

__all__ = ['SegmentationReportBuilder']
