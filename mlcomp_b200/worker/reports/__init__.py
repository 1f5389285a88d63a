from .classification import ClassificationReportBuilder
from .segmentation import SegmentationReportBuilder

__all__ = ['ClassificationReportBuilder', 'SegmentationReportBuilder']
