from .loss import FocalLossWithLogits, LabelSmoothingLossWithLogits, WeightedLoss  # noqa: F401
from .model import BertForQuestionAnswering  # noqa: F401
